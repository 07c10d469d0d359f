#!/usr/bin/env python
"""bench.py — Mrays/s of the per-pixel ray-trace pass on B200 (BASELINE.json metric).

Workload (config["workload"]): 1024 random spheres (seed 1234), 3840x2160, max_depth 1, 1 sample — one `update`
dispatch per frame, rays == pixels.  A step renders one frame per GPU (frames of a camera batch sharded over the ranks,
scene replicated: weak scaling); with N > 1 every frame's RGBA8 image lands in rank 0's device memory (fused peer stores
over NVLink out of the shading pass, completion flags behind them).

  python bench.py [--gpus N --steps K --warmup W]          our arm (CUDA kernels through the C ABI)
  python bench.py --impl reference [...]                   the CPU restatement of the shader (oracle) on the box's
                                                           host cores, same metric / config

One JSON line on stdout (rank 0).  Every number in it is measured in this run, except the block `executed.from_capture`,
which is loaded from a committed ncu capture (profiles/*.json) and carries the commit it was taken at.
Keys beyond the base contract:
  roofline      contract model (18 flop per (ray, sphere) pair over all N spheres) against the FP32-FMA peak measured in
                this process; `model` says that the shipped kernel does not execute that work
  executed      what the shipped kernel does execute: exact tests per ray (device counters), and — from the committed
                capture — warp instructions per launch, issue-slot fraction, executed FP32 flop against the FMA peak
  bruteforce    the design the port target describes (every sphere tested for every ray, shared-memory tiles, FFMA2),
                timed in the same run: the arm for which a fraction of the FP32 peak is a meaningful figure
  fast          the opt-in tolerance mode (rt_set_precision): time and mismatch count against the oracle
  strong        ONE 7680x4320 image split by row blocks over the N ranks, gathered into rank 0 (BASELINE config 4)
  weak_like_for_like   the headline step with the RGBA8 delivery every N > 1 rank performs, also at N = 1
  sustained     the headline update looped for >= 2 s with NVML clock / power samples
  parity, parity_libm, deep, hbm, cpu_baseline, e2e, clocks, gpu_launches
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W, H, N_SPHERES = 3840, 2160, 1024
W8K, H8K = 7680, 4320
FLOP_PER_TEST = 18   # oc 3 + h 5 + c 7 + disc 3, a = d.d hoisted (the reference's test as written)
METRIC = "Mrays/s at 4K/1024 spheres"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# stdout carries exactly ONE line, the JSON: everything else that libraries write to fd 1 (e.g. NCCL's
# "NCCL version ..." banner) is sent to stderr by pointing fd 1 at fd 2 for the duration of the run.
_JSON_FD = None


def capture_stdout():
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit_json(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    os.write(_JSON_FD if _JSON_FD is not None else 1, data)


def host_threads():
    """Cores this process may use.  torchrun exports OMP_NUM_THREADS=1, so omp_get_max_threads() is not the answer."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def workload_config(n_gpus):
    return {
        "workload": f"{N_SPHERES} random spheres (seed 1234) at {W}x{H}, max_depth 1, 1 sample/pixel, one frame per GPU per step",
        "spheres": N_SPHERES, "width": W, "height": H, "max_depth": 1, "frames_per_step": n_gpus,
        "parallelism": f"frames sharded over {n_gpus} GPU(s), scene replicated" + (
            ", every frame's RGBA8 image delivered into rank 0's device memory over NVLink (CUDA-IPC mapping): written locally "
            "by the shading pass and sent by a copy engine behind the next frame (`copy`; the other way, stores out of the "
            "shading pass with a release flag behind, is timed beside it); weak_like_for_like.mode says which the value used" if n_gpus > 1 else ""),
        "l2": "accumulation image (132.7 MB float4, stored once per step) exceeds the 126 MB L2; the 32 KB sphere buffer is shared-memory resident by design",
    }


class NvmlSampler:
    """SM clock / power / throttle reasons of one GPU, polled through NVML from a thread while the GPU works
    (`nvidia-smi -lms` does not produce a sample inside a few-millisecond region)."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, local_rank, period_s=0.0008):
        self.samples, self.period, self.ok, self.err = [], period_s, False, None
        self._stop = threading.Event()
        self.thread = None
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            pr = torch.cuda.get_device_properties(local_rank)
            try:
                bdf = f"{pr.pci_domain_id:08x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
                self.h = pynvml.nvmlDeviceGetHandleByPciBusId(bdf.encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
            self.nv = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:   # no NVML: say so in the record
            self.err = repr(e)

    def start(self):
        if self.ok:
            self.thread = threading.Thread(target=self._run, daemon=True)
            self.thread.start()

    def _run(self):
        nv, h = self.nv, self.h
        while not self._stop.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                mw = nv.nvmlDeviceGetPowerUsage(h)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.samples.append((time.perf_counter(), float(mhz), mw / 1000.0, int(rs)))
            except Exception as e:
                self.err = repr(e)
                break
            time.sleep(self.period)

    def stop(self):
        self._stop.set()
        if self.thread:
            self.thread.join(timeout=2)

    def window(self, t0, t1):
        """Summary of the samples taken in [t0, t1] (perf_counter): the clocks line of the JSON record."""
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: " + str(self.err)], "samples": 0}
        sel = [s for s in self.samples if t0 <= s[0] <= t1]
        bits = 0
        for s in sel:
            bits |= s[3]
        return {"sm_mhz": statistics.median(s[1] for s in sel) if sel else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(name for bit, name in self.REASONS.items() if bits & bit),
                "power_w_max": max((s[2] for s in sel), default=None), "samples": len(sel), "source": "nvml thread"}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def ffma_peak_tflops(device):
    """FP32 FMA peak measured now, on this GPU (tools/ffma_peak.cu)."""
    lib = C.CDLL(os.path.join(ROOT, "tools", "libffma_peak.so"))
    tf, sms = C.c_double(0), C.c_int(0)
    rc = lib.ffma_peak(device, 8, C.byref(tf), None, C.byref(sms))
    if rc != 0:
        raise RuntimeError(f"ffma_peak failed: {rc}")
    return tf.value, sms.value


def git_head():
    """The commit the library was built at: from git where the checkout has its history, else from the note
    __graft_entry__.build() leaves beside the library (a gpurun snapshot travels without .git)."""
    try:
        head = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
        if head:
            return head
    except Exception:
        pass
    try:
        with open(os.path.join(ROOT, "gpu_ray_tracing_b200", "librt_b200.so.commit")) as f:
            return f.read().strip() or None
    except OSError:
        return None


def committed_capture():
    """The ncu counters of the dominant kernel, from the newest committed capture (profiles/r*_primary_capture.json).
    NOT measured in this run: the block says which commit and workload it was taken at."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_primary_capture.json")))
    if not files:
        return None
    try:
        with open(files[-1]) as f:
            cap = json.load(f)
        cap["file"] = os.path.relpath(files[-1], ROOT)
        return cap
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------
# CPU arms (the only places the oracle is executed outside tests/): inputs from committed bytes + the oracle's own
# camera derivation (tests/ref_inputs.py), so these legs never map the product library.
# ------------------------------------------------------------------------------------------------
def cpu_frame(cam, rows, row_begin, threads, scene=None, libm=False):
    """One oracle `update` over a row window of the 4K workload.  Returns (accum, rays, seconds)."""
    import oracle_lib as ol
    import ref_inputs as ri
    scene = scene or ri.random_scene_1024()
    t0 = time.perf_counter()
    accum, rays = ol.oracle_update(cam, scene, W, H, row_begin=row_begin, row_count=rows, threads=threads, libm=libm)
    return accum, rays, time.perf_counter() - t0


def run_reference(args):
    """--impl reference: the reference's shader math on the host cores.  The WGSL itself cannot run here (no wgpu /
    graphics stack / Rust: 'unavailable on this box'), so this is the C++ port — every host core, whatever
    OMP_NUM_THREADS says (torchrun sets it to 1)."""
    import ref_inputs as ri
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    rows = 216   # bounded sample: a tenth of the frame
    row_begin = (H - rows) // 2
    scene = ri.random_scene_1024()
    for _ in range(max(args.warmup, 1)):
        cpu_frame(ri.camera(W, H), 48, row_begin, threads, scene)
    t0 = time.perf_counter()
    rays_total = 0
    for i in range(args.steps):
        _, rays, _ = cpu_frame(ri.camera(W, H, random_seed=(i * 2654435761 % (1 << 24)) / float(1 << 24)), rows, row_begin, threads, scene)
        rays_total += rays
    dt = time.perf_counter() - t0
    value = rays_total / dt / 1e6
    sample = (f"rows {row_begin}..{row_begin + rows} of the {W}x{H} frame per step ({rows * W} rays x {N_SPHERES} spheres): the centred tenth "
              "of the frame, where the spheres are (at depth 1 every ray tests every sphere, so the band's cost per ray is the frame's)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "CPU restatement of compute_shader.wgsl (oracle/rt_oracle.cpp, g++ -O2 -ffp-contract=off, OpenMP over rows); "
                "the reference shader via wgpu/lavapipe is unavailable on this box.  Inputs: tests/golden/random1024_seed1234.bin "
                "and the oracle's own SceneCamera::from",
    }
    emit_json(line)


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
class Gather:
    """Rank 0's gather buffer, mapped by every rank: [2 slots x n_ranks x frame_bytes | flags].  flags: arrived[rank]
    (u32, frame counter released behind that rank's stores) and one `consumed` word (rank 0: frames it has seen whole)."""

    def __init__(self, rt, dist, torch, rank, world, local_rank, slot_bytes):
        self.rank, self.world, self.slot_bytes = rank, world, slot_bytes
        self.flags_off = 2 * slot_bytes
        total = self.flags_off + 4096
        handle = torch.zeros(64, dtype=torch.uint8, device="cuda")
        if rank == 0:
            self.base = rt.device_alloc(local_rank, total)
            self.zero(torch)
            handle.copy_(torch.frombuffer(bytearray(rt.ipc_export(local_rank, self.base)), dtype=torch.uint8))
        if world > 1:
            dist.broadcast(handle, src=0)
            if rank != 0:
                self.base = rt.ipc_open(local_rank, bytes(handle.cpu().numpy().tobytes()))

    def zero(self, torch):
        from cuda.bindings import runtime as cudart
        err, = cudart.cudaMemset(self.base + self.flags_off, 0, 4096)
        assert int(err) == 0, err
        torch.cuda.synchronize()

    def slot(self, step):
        return self.base + (step & 1) * self.slot_bytes

    def arrived(self, of_rank=0):
        return self.base + self.flags_off + 4 * of_rank

    def consumed(self):
        return self.base + self.flags_off + 2048


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import gpu_ray_tracing_b200 as rt
    import scenes

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        log(f"note: WORLD_SIZE={world} but --gpus {args.gpus}; using WORLD_SIZE")
    if rt.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device (no CPU fallback): " + rt.lib().rt_last_error().decode())
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()

    def allreduce(vals, op):
        if world == 1:
            return [float(x) for x in vals]
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return [float(x) for x in t.tolist()]

    K, Wm = args.steps, max(args.warmup, 3)
    scene = scenes.random_scene(N_SPHERES)

    # one camera per (step, rank): the same pose with a re-rolled random_seed (camera.rs:346) — depth-1
    # cost is pose independent (every ray tests every sphere), so steps are comparable
    def cam_for(step, of_rank=None, depth=1, spp=1, w=W, h=H):
        k = rank if of_rank is None else of_rank
        seed = ((step * world + k) * 2654435761 % (1 << 24)) / float(1 << 24)
        return scenes.camera(w, h, depth, spp, random_seed=seed)

    r = rt.Renderer(W, H, device=local_rank)
    r.set_scene(scene)
    r.init()
    frame_bytes = W * H * 4
    sampler = NvmlSampler(local_rank)
    sampler.start()

    # ---- gather target: rank 0 owns 2 slots x world RGBA8 frames + flags; every rank's shading pass stores its frame's
    # RGBA8 texels straight into its place there (rt_set_rgba8_target: local stores on rank 0, NVLink stores elsewhere) and
    # releases arrived[rank] = step + 1 behind them; rank 0 waits for the flags of the last step before the clock stops.
    gather = Gather(rt, dist, torch, rank, world, local_rank, world * frame_bytes) if (world > 1 or args.like_for_like) else None

    # Two ways to deliver a frame (args.deliver):
    #   fused  the shading pass stores the RGBA8 texels into the slot itself (rt_set_rgba8_target) and a release flag follows;
    #          nothing runs after the kernel, but the stores cross NVLink while the SMs compute
    #   copy   the shading pass writes the frame into a local buffer (rt_set_rgba8_output) and a copy engine sends it on a copy
    #          stream while the next frame renders (rt_read_rgba8_async): decoupled from the SMs, one frame of latency
    # auto = copy: measured at every N (profiles/r2_bench_n*_final.json, both ways timed in each run) the copy engines
    # deliver a frame at less cost to the step than the shading pass's own stores — N = 2 / 4: 0.277 against 0.287 ms per
    # step, N = 8 (rank 0's ingress at the link's bandwidth, where the stores back-pressure the SMs that issue them): 0.296
    # against 0.356.  Fused delivery remains what a caller picks for latency: its frame is whole when the kernel ends.
    mode = args.deliver if args.deliver != "auto" else "copy"

    def weak_step(i, how):
        if how == "fused":
            r.set_rgba8_target(gather.slot(i) + rank * frame_bytes)
        r.update(cam_for(i))
        if how == "fused":
            r.stream_signal(gather.arrived(rank), i + 1)
        elif how == "copy":
            r.read_rgba8_async_into(gather.slot(i) + rank * frame_bytes)

    def weak_region(n_steps, first, how):
        """n_steps headline steps, device-timed on the renderer's stream; returns (ms, wall window)."""
        r.set_rgba8_target(None)
        r.set_rgba8_output(how == "copy")
        for i in range(3):
            weak_step(i, how)
        if how == "copy":
            r.read_wait()
        r.synchronize()
        torch.cuda.synchronize()
        r.reset_counters()
        l0 = r.kernel_launches()
        barrier()
        t0 = time.perf_counter()
        r.timer_begin()
        for i in range(n_steps):
            weak_step(first + i, how)
        if how == "fused" and rank == 0:
            r.stream_wait(gather.arrived(0), world, first + n_steps, 5000)   # every rank's last frame is whole
        if how == "copy":
            r.read_wait()                                                   # ... has left this rank (host wait, then the end event)
        ms = r.timer_end()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        barrier()
        r.set_rgba8_target(None)
        r.set_rgba8_output(False)
        return ms, (t0, t1), r.rays_traced(), r.kernel_launches() - l0

    how = mode if world > 1 else "none"
    for i in range(Wm):
        weak_step(i, "none")
    r.synchronize()
    ms, region, rays, launches = weak_region(K, Wm, how)
    timed_out = r.stream_wait_timed_out()
    mx = allreduce([ms], dist.ReduceOp.MAX)
    sm = allreduce([float(rays), float(launches)], dist.ReduceOp.SUM)
    ms, rays, launches = mx[0], int(sm[0]), int(sm[1])
    value = rays / (ms * 1e-3) / 1e6
    clocks = sampler.window(*region) if rank == 0 else None

    # ---- like-for-like basis for the scaling efficiency: the N = 1 headline above skips the RGBA8 delivery because there is
    # nothing to gather; here the same step WITH it — at N = 1 both ways (into a local buffer), at N > 1 the way not used above
    like = None
    if args.like_for_like:
        def timed(how_):
            ms_ = weak_region(K, Wm, how_)[0]
            ms_ = allreduce([ms_], dist.ReduceOp.MAX)[0]
            return {"ms_per_step": ms_ / K, "value": world * K * W * H / (ms_ * 1e-3) / 1e6}
        if world == 1:
            like = {"unit": "Mrays/s", "fused": timed("fused"), "copy": timed("copy"),
                    "what": "the headline step + the delivery of the frame's RGBA8 image every rank performs at N > 1 (into a local "
                            "buffer here): `fused` = stored by the shading pass + completion flag, `copy` = written locally and sent "
                            "by a copy engine behind the next frame.  Efficiency at N compares value(N) with N x the figure of the "
                            "mode that run reports in weak_like_for_like.mode"}
        else:
            other = "copy" if mode == "fused" else "fused"
            like = {"unit": "Mrays/s", "mode": mode, mode: {"ms_per_step": ms / K, "value": value}, other: timed(other),
                    "rank0_ingress_gbs": (world - 1) * frame_bytes * K / (ms * 1e-3) / 1e9, "nvlink_ingress_peak_gbs": 770.0,
                    "what": "same work per GPU as weak_like_for_like[mode] of the N = 1 run; rank 0 takes in (N-1) x 33.2 MB per step over "
                            "NVLink (770 GB/s per direction measured for a peer copy on this pool, 900 nominal): gather-to-one-GPU is "
                            "the ceiling of this leg"}

    # ---- the gathered frames are the frames: rank 0 re-renders rank 1's last frame and compares it with its slot ---
    gather_check = None
    if world > 1:
        barrier()
        if rank == 0:
            last = Wm + K - 1
            r.update(cam_for(last, of_rank=1))
            expect = r.read_rgba8()
            got = device_bytes(torch, gather.slot(last) + 1 * frame_bytes, frame_bytes).reshape(H, W, 4)
            gather_check = {"frame": "last step of rank 1, re-rendered on rank 0", "identical": bool(np.array_equal(got, expect)),
                            "waits_timed_out": timed_out}
            r.init()
        barrier()

    # ---- strong scaling: ONE 7680x4320 frame split by row blocks over the ranks (BASELINE config 4) ----------------
    strong = strong_leg(rt, dist, torch, np, scenes, scene, rank, world, local_rank, barrier, allreduce, args)

    # ---- dominant kernel alone (roofline): CUDA events on the launching stream ------------------
    n_k = 20
    cam = cam_for(0)
    r.update(cam)
    r.synchronize()
    r.reset_counters()
    r.timer_begin()
    for _ in range(n_k):
        r.update(cam)
    kernel_ms = r.timer_end() / n_k
    exact_per_ray = r.exact_tests() / max(1, r.rays_traced())
    flops_per_launch = float(W) * H * N_SPHERES * FLOP_PER_TEST
    achieved_tf = flops_per_launch / (kernel_ms * 1e-3) / 1e12

    if rank != 0:
        # e2e is measured on every rank at once (N host read-backs share the host side)
        e2e = e2e_leg(rt, torch, np, r, scene, cam_for, local_rank, world, K, barrier, allreduce)
        barrier()
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- brute force (variant 2): every sphere for every ray from shared-memory tiles, FFMA2 broad phase -----------
    r.set_kernel_variant(2)
    r.update(cam)
    r.synchronize()
    r.timer_begin()
    for _ in range(5):
        r.update(cam)
    bf_ms = r.timer_end() / 5
    bf_img = r.read_accum()
    r.set_kernel_variant(0)

    # ---- tolerance modes -----------------------------------------------------------------------
    fast_runs = {}
    for name, level in (("fast_shading", rt.RT_PRECISION_FAST_SHADING), ("fast_all", rt.RT_PRECISION_FAST_ALL)):
        r.set_precision(level)
        r.update(cam)
        r.synchronize()
        r.timer_begin()
        for _ in range(n_k):
            r.update(cam)
        fast_runs[name] = (r.timer_end() / n_k, r.read_accum())
    r.set_precision(rt.RT_PRECISION_EXACT)
    r.update(cam)
    exact_img = r.read_accum()

    # ---- the reference's default depth (max_depth 30), rays counted by the kernels -------
    deep, deep_img, deep_fast_img, cam30 = deep_leg(rt, np, scenes, r, scene, local_rank)

    # ---- sustained: the headline update looped for >= 2 s, clocks and power sampled through NVML ------------------
    sustained = sustained_leg(r, cam_for, sampler, args.sustain_s if world == 1 else 0.0)
    if clocks is not None and clocks.get("samples", 0) < 3 and sustained and sustained.get("samples", 0) >= 3:
        # a timed region of a few milliseconds can fall between two NVML polls: the same update under sustained load
        clocks = {k: sustained[k] for k in ("sm_mhz", "sm_max_mhz", "reasons", "power_w_max", "samples")}
        clocks["source"] = "nvml thread during the sustained leg (the timed region held fewer than 3 samples)"

    # ---- end to end through the C ABI with HOST buffers ----------------------------------------
    e2e = e2e_leg(rt, torch, np, r, scene, cam_for, local_rank, world, K, barrier, allreduce)
    sampler.stop()

    # ---- denominators, CPU baseline, parity, the JSON line ---------------------------------------
    import oracle_lib as ol
    import ref_inputs as ri
    peak_tf, sms = ffma_peak_tflops(local_rank)
    peaks = measured_peaks()
    sm_max = (clocks or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
    theo_tf = sms * 128 * 2 * sm_max * 1e6 / 1e12
    hbm_bytes = float(W) * H * 16      # reset frames: one float4 store per pixel, nothing read
    hbm_gbs = hbm_bytes / (kernel_ms * 1e-3) / 1e9
    threads = host_threads()
    cpu, parity, parity_libm = None, None, None
    ref_scene = ri.random_scene_1024()
    same_inputs = bool(np.array_equal(ref_scene.spheres.view(np.uint8), scene.spheres.view(np.uint8)))
    if world == 1 and not args.no_cpu_baseline:
        # bounded sample: whole frames of the same workload until ~10 s of wall time (>= 2 frames); the FIRST of them
        # is the frame the GPU rendered above and is compared texel by texel — all 8,294,400 of them
        cpu_rays, dt, frames = 0, 0.0, 0
        while frames < 2 or (dt < 10.0 and frames < 8):
            img, r1, t1 = cpu_frame(cam if frames == 0 else cam_for(frames), H, 0, threads, ref_scene)
            if frames == 0:
                parity = ol.compare_images(img, exact_img)
                parity["checked"] = "the whole benchmark frame (camera of step 0) vs oracle/rt_oracle.cpp, same bytes in"
                parity["inputs_identical_to_fixture"] = same_inputs
                parity["oracle"] = "parity unpinned by the reference (no vectors, WGSL cannot run here); sin/cos/pow5 from include/rt_pinned_math.h on both sides"
                fast_rep = {k: ol.compare_images(img, v[1]) for k, v in fast_runs.items()}
                bf_rep = ol.compare_images(img, bf_img)
            cpu_rays += r1
            dt += t1
            frames += 1
        cpu = {"value": cpu_rays / dt / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "port",
               "sample": f"{frames} full frames of the same workload ({cpu_rays} rays x {N_SPHERES} spheres) in {dt:.1f} s, "
                         "oracle/rt_oracle.cpp (g++ -O2 -ffp-contract=off) with OpenMP over rows"}
        # the mismatch count the port target asks for: against an implementation with DIFFERENT builtin rounding (libm
        # sinf / cosf / powf instead of the pinned polynomials), at depth 1 and at the reference's depth 30
        libm1, _, _ = cpu_frame(cam, H, 0, threads, ref_scene, libm=True)
        rep1 = ol.compare_images(libm1, exact_img)
        libm30, _ = ol.oracle_update(cam30, ref_scene, W, H, threads=threads, libm=True)
        pin30, _ = ol.oracle_update(cam30, ref_scene, W, H, threads=threads)
        rep30 = ol.compare_images(libm30, deep_img)
        parity_libm = {
            "what": "GPU (exact mode) vs liboracle_libm.so: the same restatement with libm sinf/cosf/powf — a stand-in for a WGSL "
                    "implementation whose builtins round differently; tolerance 1/255 per channel after RGBA8 quantisation",
            "depth1": {k: rep1[k] for k in ("pixels", "pixels_over_1_255", "pixels_differing", "max_abs")},
            "depth30": {k: rep30[k] for k in ("pixels", "pixels_over_1_255", "pixels_differing", "max_abs")},
            "depth30_pinned_vs_libm_oracles": {k: v for k, v in ol.compare_images(libm30, pin30).items() if k != "bit_identical"},
        }
        deep["parity"] = ol.compare_images(pin30, deep_img)
        if deep_fast_img is not None:
            rep = ol.compare_images(pin30, deep_fast_img)
            deep["fast"]["vs_oracle"] = {k: rep[k] for k in ("pixels", "pixels_over_1_255", "pixels_differing", "max_abs")}
    else:
        # N > 1 (or --no-cpu-baseline): a band of the frame against the oracle, the arms against the exact image
        rows, row_begin = 216, (H - 216) // 2
        img, _, _ = cpu_frame(cam, rows, row_begin, threads, ref_scene)
        parity = ol.compare_images(img, exact_img[row_begin:row_begin + rows])
        parity["checked"] = f"rows {row_begin}..{row_begin + rows} of the benchmark frame vs oracle (the whole frame is compared in the N = 1 run)"
        fast_rep = {k: ol.compare_images(exact_img, v[1]) for k, v in fast_runs.items()}
        bf_rep = ol.compare_images(exact_img, bf_img)

    capture = committed_capture()
    executed = {
        "exact_tests_per_ray": exact_per_ray,
        "what": "device counters of this run: (ray, sphere) pairs that reach the reference quadratic after the tile / tree / sign-test "
                "cascade; the other N - that many pairs are discarded by conservative tests with identical results",
        "from_capture": capture,
    }
    if capture and capture.get("executed_fp32_flop_per_launch"):
        executed["fp32_tflops"] = capture["executed_fp32_flop_per_launch"] / (kernel_ms * 1e-3) / 1e12
        executed["fp32_frac_of_peak"] = executed["fp32_tflops"] / peak_tf
        executed["fp32_note"] = "captured FADD + FMUL + 2 x FFMA thread instructions / kernel_ms of this run / FFMA peak of this run: the kernel is issue-bound, not FMA-bound"
    if capture and capture.get("warp_instructions_per_launch") and clocks and clocks.get("sm_mhz"):
        slots = sms * 4 * clocks["sm_mhz"] * 1e6 * kernel_ms * 1e-3
        executed["issue_slot_frac"] = capture["warp_instructions_per_launch"] / slots
        executed["issue_slot_frac_note"] = "captured warp instructions / (SMs x 4 schedulers x median SM clock of this run x kernel_ms of this run)"
    roofline_executed = {
        "executed_bound": "issue slots (4 warp schedulers per SM, one instruction per cycle each)",
        "executed_frac": executed.get("issue_slot_frac"),
        "executed_fp32_frac_of_peak": executed.get("fp32_frac_of_peak"),
        "executed_note": "the efficiency figure of the shipped kernel: captured warp instructions / issue slots of this run's launch; "
                         "`frac` above follows the contract's brute-force model and is not an efficiency",
    }
    line = {
        "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(world),
        "roofline": {"bound": "fp32_fma", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": achieved_tf / peak_tf, "traffic": (capture or {}).get("dram_bytes_per_launch"),
                     "traffic_source": (capture or {}).get("file"),
                     "model": "18 flop x N spheres per ray (SURVEY.md 8(d)) — brute force, NOT executed by this kernel: frac > 1 says "
                              "the kernel reaches the same nearest hits without that work, it is not an efficiency. See `executed` "
                              "for what it does execute and `bruteforce` for the arm that does the modelled work",
                     "peak_source": "FFMA microbenchmark run in this process (tools/ffma_peak.cu); MEASURED_PEAKS.json has no FP32 figure",
                     "frac_of_theoretical": achieved_tf / theo_tf, "theoretical_peak": theo_tf,
                     "algorithmic_flop_per_launch": flops_per_launch, "kernel_ms": kernel_ms,
                     "kernel": "primary_hits_kernel<kPrimShade, exact> (+ prep_primary_tree_kernel, ~4 us)", **roofline_executed},
        "executed": executed,
        "bruteforce": {"kernel": "update_resident_kernel (variant 2): all 1024 records from shared memory for every ray, 7-FFMA2 "
                                 "conservative test per pair + the reference quadratic on its candidates",
                       "ms": bf_ms, "mrays_s": W * H / bf_ms / 1e3, "tflops": flops_per_launch / (bf_ms * 1e-3) / 1e12,
                       "frac": flops_per_launch / (bf_ms * 1e-3) / 1e12 / peak_tf,
                       "bit_identical_to_oracle": bf_rep["bit_identical"]},
        "fast": {"note": "opt-in tolerance modes of rt_set_precision (the default and the headline are exact): FAST_SHADING = MUFU + FMA "
                         "in scatter / sky / running mean with exact primary rays; FAST_ALL = ray generation too (silhouette samples may flip)",
                 **{name: {"kernel_ms": t, "mrays_s": W * H / t / 1e3, "speedup_vs_exact": kernel_ms / t,
                           "vs_oracle": {k: fast_rep[name][k] for k in ("pixels", "pixels_over_1_255", "pixels_differing", "max_abs")}}
                    for name, (t, _) in fast_runs.items()}},
        "hbm": {"achieved": hbm_gbs, "peak": peaks.get("hbm_gbs"), "unit": "GB/s",
                "frac": (hbm_gbs / peaks["hbm_gbs"]) if peaks.get("hbm_gbs") else None,
                "algorithmic_bytes_per_launch": hbm_bytes,
                "note": "one float4 accumulator store per pixel (camera_has_moved = 1: no read); not the bound"},
        "cpu_baseline": cpu, "e2e": e2e, "deep": deep, "strong": strong, "weak_like_for_like": like,
        "sustained": sustained, "gather_check": gather_check, "gpu_launches": int(launches), "clocks": clocks,
        "parity": parity, "parity_libm": parity_libm,
        "tree": tree_block(rt, scenes, r, local_rank),
        "git_head": git_head(),
    }
    emit_json(line)
    barrier()
    if world > 1:
        dist.destroy_process_group()


def device_bytes(torch, ptr, nbytes):
    """Copy `nbytes` at device address `ptr` (current device) to a numpy array."""
    from cuda.bindings import runtime as cudart
    host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    err, = cudart.cudaMemcpy(host.data_ptr(), ptr, nbytes, cudart.cudaMemcpyKind.cudaMemcpyDeviceToHost)
    assert int(err) == 0, err
    return host.numpy()


def strong_leg(rt, dist, torch, np, scenes, scene, rank, world, local_rank, barrier, allreduce, args):
    """One 7680x4320 / 1024-sphere frame per step, dealt over the ranks as interleaved 16-row bands (rt_create_bands), the
    RGBA8 rows gathered into rank 0's full frame over NVLink.  Two gathers are timed:
      pipelined  the shading pass writes the rows locally, a copy engine scatters them into rank 0's frame on the copy
                 stream while the next frame renders (rt_send_rgba8_async): the link carries frame f behind the kernels of
                 f + 1 — the `ms` of this leg
      fused      the shading pass stores the rows straight into rank 0's frame (rt_set_rgba8_target): no frame of latency,
                 but the stores share NVLink with the kernel that issues them
    Protocol per frame f and rank i, both ways: wait consumed >= f - 1 (the slot's previous frame has been seen whole) ->
    deliver -> release arrived[i] = f + 1; rank 0 waits for every arrived[*] >= f + 1 — the frame is whole in its memory —
    and releases consumed = f + 1.  Time = device time, max over ranks, for K whole frames.  Afterwards rank 0 renders the
    same frames alone (t1) and compares."""
    # frames per timed leg, whatever --steps says (that is the headline's K): the pipelined gather ends on a drain of one
    # frame's copies, which the leg's time includes — over 20 frames that is 5 % of the figure, over 100 it is 1 %
    K = 100
    frame8 = W8K * H8K * 4
    band_rows = int(os.environ.get("RT_BENCH_BAND_ROWS", "16"))   # (a multiple of 8; A/B knob)
    g = Gather(rt, dist, torch, rank, world, local_rank, frame8)
    # interleaved 16-row bands, not contiguous blocks: rows that hold geometry cost more than sky rows
    rs = rt.Renderer(W8K, H8K, device=local_rank, bands=(band_rows, world, rank))
    rs.set_scene(scene)
    rs.init()

    def cam8(f):
        return scenes.camera(W8K, H8K, 1, 1, random_seed=((f + 1) * 2654435761 % (1 << 24)) / float(1 << 24))

    def frame_fused(f):
        if f >= 2:
            rs.stream_wait(g.consumed(), 1, f - 1, 5000)
        rs.set_rgba8_target(g.slot(f))
        rs.update(cam8(f))
        rs.stream_signal(g.arrived(rank), f + 1)
        if rank == 0:
            rs.stream_wait(g.arrived(0), world, f + 1, 5000)
            rs.stream_signal(g.consumed(), f + 1)

    def frame_pipelined(f):
        if rank == 0:
            # the owner's rows need no link: stored in place by the shading pass; its copy stream only keeps the books
            if f >= 2:
                rs.stream_wait(g.consumed(), 1, f - 1, 5000)
            rs.set_rgba8_target(g.slot(f))
            rs.update(cam8(f))
            rs.stream_signal(g.arrived(0), f + 1)
            rs.copy_wait(g.arrived(0), world, f + 1, 5000)
            rs.copy_signal(g.consumed(), f + 1)
            return
        rs.update(cam8(f))                                   # (rt_set_rgba8_output: the rows land in a local buffer)
        if f >= 2:
            rs.copy_wait(g.consumed(), 1, f - 1, 5000)
        rs.send_rgba8_async(g.slot(f))
        rs.copy_signal(g.arrived(rank), f + 1)

    def timed(frame_fn, first, pipelined):
        rs.set_rgba8_target(None)
        rs.set_rgba8_output(pipelined and rank != 0)
        for f in range(first, first + 3):
            frame_fn(f)
        rs.read_wait()
        rs.synchronize()
        torch.cuda.synchronize()
        barrier()
        rs.timer_begin()
        for f in range(first + 3, first + 3 + K):
            frame_fn(f)
        if pipelined:
            rs.read_wait()                                   # the copy stream has delivered (and rank 0 has seen) the last frame
        ms_ = rs.timer_end()
        torch.cuda.synchronize()
        barrier()
        got_ = device_bytes(torch, g.slot(first + 3 + K - 1), frame8).reshape(H8K, W8K, 4).copy() if rank == 0 else None
        barrier()
        return ms_, got_, first + 3 + K - 1

    ms_p, got_p, last_p = timed(frame_pipelined, 0, True)
    ms_f, got_f, last_f = timed(frame_fused, 3 + K, False)
    timed_out = rs.stream_wait_timed_out()
    # without the gather: the rows stay where they are rendered
    rs.set_rgba8_target(None)
    rs.set_rgba8_output(False)
    rs.update(cam8(0))
    rs.synchronize()
    barrier()
    rs.timer_begin()
    for f in range(K):
        rs.update(cam8(f))
    ms_nog = rs.timer_end()
    barrier()
    mx = allreduce([ms_p, ms_f, ms_nog, float(timed_out)], dist.ReduceOp.MAX)
    ms_p, ms_f, ms_nog, timed_out = mx[0] / K, mx[1] / K, mx[2] / K, int(mx[3])
    rs.close()
    barrier()
    out = None
    if rank == 0:
        # the same frames on this GPU alone: t1, and the images the gathered ones must equal
        with rt.Renderer(W8K, H8K, device=local_rank) as r1:
            r1.set_scene(scene)
            r1.init()
            r1.set_rgba8_target(g.slot(0))
            r1.update(cam8(last_p))
            r1.synchronize()
            same_p = bool(np.array_equal(got_p, device_bytes(torch, g.slot(0), frame8).reshape(H8K, W8K, 4)))
            r1.update(cam8(last_f))
            r1.synchronize()
            same_f = bool(np.array_equal(got_f, device_bytes(torch, g.slot(0), frame8).reshape(H8K, W8K, 4)))
            r1.timer_begin()
            for f in range(K):
                r1.update(cam8(last_f))
            t1 = r1.timer_end() / K
        inbound = (world - 1) / world * frame8
        out = {"workload": f"one {W8K}x{H8K} frame / {N_SPHERES} spheres / depth 1 per step, interleaved {band_rows}-row bands dealt over "
                           f"{world} rank(s) (rt_create_bands), RGBA8 rows gathered into rank 0's frame over NVLink, arrived / consumed "
                           "flags per frame; `ms` = the pipelined gather (copy engines behind the next frame), `fused` = stores out of "
                           "the shading pass",
               "ms": ms_p, "mrays_s": W8K * H8K / ms_p / 1e3, "t1_ms_same_run": t1, "speedup_vs_1": t1 / ms_p,
               "rank0_ingress_gbs": inbound / (ms_p * 1e-3) / 1e9 if world > 1 else 0.0, "nvlink_ingress_peak_gbs": 770.0,
               "fused": {"ms": ms_f, "speedup_vs_1": t1 / ms_f, "rank0_ingress_gbs": inbound / (ms_f * 1e-3) / 1e9 if world > 1 else 0.0,
                         "identical_to_single_gpu_frame": same_f},
               "ms_without_gather": ms_nog, "speedup_vs_1_without_gather": t1 / ms_nog,
               "link_floor_ms": inbound / 770e9 * 1e3,
               "identical_to_single_gpu_frame": same_p, "frames": K, "waits_timed_out": timed_out}
    barrier()
    return out


def tree_block(rt, scenes, r, local_rank):
    """Host time of the tree build (rt_tree.h: built on one host thread at the first update after new sphere bytes) for the
    benchmark scene and for the 65,536-sphere scene of config 5, next to rt_tree_info's (nodes, levels, bounce path)."""
    out = {"build_ms_1024_spheres": r.tree_build_ms(), "info": r.tree_info()}
    with rt.Renderer(64, 36, device=local_rank) as rb:
        rb.set_scene(scenes.grid_scene(65536))
        rb.init()
        rb.update(rt.orbit_camera(3, 256, 64, 36))
        rb.synchronize()
        out["build_ms_65536_spheres"] = rb.tree_build_ms()
        out["info_65536_spheres"] = rb.tree_info()
    return out


def deep_leg(rt, np, scenes, r, scene, local_rank):
    cam30 = scenes.camera(W, H, 30, 500)
    r.init()
    r.update(cam30)
    r.synchronize()
    r.reset_counters()
    l0 = r.kernel_launches()
    r.timer_begin()
    n_deep = 5
    for _ in range(n_deep):
        r.update(cam30)            # camera_has_moved = 1: every frame restarts the same first sample
    deep_ms = r.timer_end() / n_deep
    deep_rays = r.rays_traced() / n_deep
    deep_launches = (r.kernel_launches() - l0) / n_deep
    deep_img = r.read_accum()
    deep = {"max_depth": 30, "ms_per_frame": deep_ms, "rays_per_frame": deep_rays, "mrays_s": deep_rays / deep_ms / 1e3,
            "launches_per_frame": deep_launches,
            "note": "primary pass (first hits resolved and survivors queued in the same kernel) + 29 wavefront bounce launches over compacted path queues"}
    r.set_precision(rt.RT_PRECISION_FAST_SHADING)
    r.update(cam30)
    r.synchronize()
    r.reset_counters()
    r.timer_begin()
    for _ in range(n_deep):
        r.update(cam30)
    fast_ms = r.timer_end() / n_deep
    fast_rays = r.rays_traced() / n_deep
    deep_fast_img = r.read_accum()
    r.set_precision(rt.RT_PRECISION_EXACT)
    deep["fast"] = {"ms_per_frame": fast_ms, "rays_per_frame": fast_rays, "speedup_vs_exact": deep_ms / fast_ms}
    r.init()
    # the application's own working point (src/lib.rs:25, src/camera.rs:33-34): default scene, 1280x720, depth 30, 500 spp
    dscene = scenes.default_scene()
    with rt.Renderer(1280, 720, device=local_rank) as rd:
        rd.set_scene(dscene)
        rd.init()
        first = scenes.camera(1280, 720, 30, 500, random_seed=0.5, moved=1)
        rd.update(first)
        rd.synchronize()
        rd.timer_begin()
        for _ in range(10):
            rd.update(first)
        frame_ms = rd.timer_end() / 10
        cams = [scenes.camera(1280, 720, 30, 500, random_seed=((i + 1) * 2654435761 % (1 << 24)) / float(1 << 24), moved=1 if i == 0 else 0)
                for i in range(500)]
        rd.init()
        t0 = time.perf_counter()
        rd.timer_begin()
        rd.update_n(cams)
        spp_ms = rd.timer_end()
        wall_ms = (time.perf_counter() - t0) * 1e3
        n_done = float(rd.read_accum()[..., 3].min())
        deep["app_default"] = {"workload": "default scene (199 spheres) 1280x720, max_depth 30", "ms_per_frame": frame_ms,
                               "time_to_500_spp_ms": spp_ms, "wall_ms": wall_ms, "samples_reached": n_done,
                               "call": "rt_update_n with 500 per-frame uniforms (random_seed re-rolled per frame, camera.rs:346)"}
    return deep, deep_img, deep_fast_img, cam30


def sustained_leg(r, cam_for, sampler, seconds):
    if seconds <= 0:
        return None
    r.init()
    t0 = time.perf_counter()
    r.timer_begin()
    n, chunk = 0, 200
    while True:
        for i in range(chunk):
            r.update(cam_for(i & 63))
        n += chunk
        r.synchronize()
        if time.perf_counter() - t0 >= seconds:
            break
    ms = r.timer_end()
    t1 = time.perf_counter()
    out = sampler.window(t0, t1)
    out.update({"seconds": ms * 1e-3, "steps": n, "ms_per_step": ms / n, "mrays_s": n * W * H / ms / 1e3,
                "what": "the headline update back to back for the stated time (synchronised every 200 steps so that the queue stays short)"})
    return out


def e2e_leg(rt, torch, np, r, scene, cam_for, local_rank, world, K, barrier, allreduce):
    """Per step: sphere buffer uploaded from pinned host memory (the reference re-creates its buffers from host memory
    every frame, src/lib.rs:151-207), the 176-byte camera uniform rides in the kernel parameters, and the frame comes
    back as RGBA8 into pinned host memory.  Pinned frames are allocated from a thread bound to the GPU's own NUMA node
    when the box says which that is."""
    import torch.distributed as dist
    with numa_local(local_rank) as numa_note:
        host_spheres = torch.from_numpy(scene.spheres.view(np.uint8).copy()).pin_memory()
        host_images = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(2)]
        for t in host_images:
            t.zero_()          # first touch on this node
    K_e2e = max(3, min(K, 50))

    def e2e_step(i):
        # frame i lands in host_images[i & 1]; the read of frame i - 2 from that buffer is complete (at most two
        # reads are in flight and the third queues behind the first on the device)
        r.set_scene_from(host_spheres.data_ptr(), len(scene.spheres), scene.count)
        r.update(cam_for(i))
        r.read_rgba8_async_into(host_images[i & 1].data_ptr())

    r.set_rgba8_output(True)                          # the RGBA8 frame comes out of the shading pass
    for i in range(2):
        e2e_step(i)
    r.read_wait()
    barrier()
    t0 = time.perf_counter()
    for i in range(K_e2e):
        e2e_step(i)
    r.read_wait()                                     # every frame of the timed region is on the host
    e2e_s = time.perf_counter() - t0
    r.set_rgba8_output(False)
    e2e_max = allreduce([e2e_s], dist.ReduceOp.MAX)[0]
    per_rank_gbs = K_e2e * W * H * 4 / e2e_s / 1e9
    sum_gbs = allreduce([per_rank_gbs], dist.ReduceOp.SUM)[0]
    e2e_value = K_e2e * world * W * H / e2e_max / 1e6
    host_image = host_images[(K_e2e - 1) & 1]
    checksum = int(host_image.view(-1)[::4099].to(torch.int64).sum().item())
    # the same loop with the blocking read (rt_read_rgba8): what a strictly frame-synchronous caller sees
    t0 = time.perf_counter()
    for i in range(K_e2e):
        r.set_scene_from(host_spheres.data_ptr(), len(scene.spheres), scene.count)
        r.update(cam_for(i))
        r.read_rgba8_into(host_image.data_ptr())
    e2e_sync_value = K_e2e * W * H / (time.perf_counter() - t0) / 1e6
    return {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": int(len(scene.spheres) * 32 + 176),
            "d2h_bytes_per_step": W * H * 4, "steps": K_e2e, "pinned_host_memory": str(numa_note),
            "d2h_gbs_this_rank": per_rank_gbs, "d2h_gbs_all_ranks": sum_gbs,
            "bound": "host read-back: 33.2 MB of RGBA8 per frame over PCIe (the kernel needs about half that time)",
            "call": "rt_set_scene + rt_update + rt_read_rgba8_async per frame (two pinned host frames, read-back of "
                    "frame i overlaps the kernel of frame i+1), rt_read_wait before the clock stops",
            "blocking_read_value": e2e_sync_value, "checksum": checksum}


class numa_local:
    """Binds the calling thread to the CPUs next to GPU `index` (sysfs local_cpulist of its PCI function) for the
    duration of the block, so that host memory allocated and first touched inside lands on that NUMA node.
    Says what it did; does nothing when sysfs or the affinity call are not available."""

    def __init__(self, index):
        self.index, self.saved, self.note = index, None, "unchanged"

    def __enter__(self):
        try:
            import torch
            pr = torch.cuda.get_device_properties(self.index)
            bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
            base = f"/sys/bus/pci/devices/{bdf}/"
            node = open(base + "numa_node").read().strip()
            cpus = set()
            for part in open(base + "local_cpulist").read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
            allowed = os.sched_getaffinity(0)
            use = cpus & allowed
            if use and use != allowed:
                self.saved = allowed
                os.sched_setaffinity(0, use)
                self.note = f"bound to {len(use)} CPUs of NUMA node {node} ({bdf})"
            else:
                self.note = f"NUMA node {node}: {'no local CPU in the cpuset' if not use else 'cpuset already local'}"
        except Exception as e:   # no sysfs, no permission: allocate where we are
            self.note = f"unavailable ({type(e).__name__})"
        return self

    def __exit__(self, *exc):
        if self.saved is not None:
            os.sched_setaffinity(0, self.saved)
        return False

    def __str__(self):
        return self.note


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sustain-s", type=float, default=2.0, help="length of the sustained leg (0 = skip)")
    ap.add_argument("--deliver", choices=["auto", "fused", "copy"], default="auto",
                    help="N > 1: how a finished frame reaches rank 0 (see run_ours)")
    ap.add_argument("--no-like-for-like", dest="like_for_like", action="store_false",
                    help="N = 1: skip the second timing of the headline step with the RGBA8 delivery of the N > 1 ranks")
    args = ap.parse_args()
    capture_stdout()
    if args.impl == "reference":
        if args.steps == 400:
            args.steps, args.warmup = 10, 1
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
