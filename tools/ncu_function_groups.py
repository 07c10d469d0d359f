"""Groups the per-line attribution of an ncu capture of primary_hits_kernel by the FUNCTION each source line belongs to
(line ranges found by pattern in the current sources), so the instruction budget can be itemised.
Usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass | python tools/ncu_line_summary.py 5000 > lines.txt
       python tools/ncu_function_groups.py lines.txt"""
import os,re,sys
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lines=open(sys.argv[1]).read().split('\n')
rows=[]
for l in lines:
    m=re.match(r"\s+(\S+):(\d+)\s+inst\s+([\d.]+)% samples\s+([\d.]+)%",l)
    if m: rows.append((m.group(1),int(m.group(2)),float(m.group(3)),float(m.group(4))))
def ranges(path, names):
    src=open(path).read().split('\n')
    out=[]
    for name,pat in names:
        for i,l in enumerate(src):
            if re.search(pat,l): out.append((i+1,name)); break
        else: print("missing",name)
    out.sort()
    return out
core=ranges(os.path.join(ROOT,'gpu_ray_tracing_b200','csrc','rt_core.cuh'),[
 ("vector helpers (add/sub/mul/dot/divs/normalize)", r"^struct f3"),
 ("fastm / templated helpers", r"^namespace fastm"),
 ("hash, random_float", r"^RT_DEV uint32_t f2u"),
 ("random_unit_vector", r"RT_DEV f3 random_unit_vector"),
 ("get_ray", r"^struct Ray"),
 ("sphere_root (exact quadratic)", r"RT_DEV bool sphere_root"),
 ("sky_times", r"RT_DEV f3 sky_times"),
 ("scatter", r"RT_DEV bool scatter_impl"),
 ("broad phase (general)", r"^#define RT_KAPPA"),
 ("primary_arm (unit direction)", r"RT_DEV bool primary_arm"),
 ("primary_h / candidate", r"RT_DEV float primary_h"),
 ("tile test (primary_tile_*)", r"RT_DEV float primary_m_cam"),
 ("primary_near / cluster", r"RT_DEV float primary_near\("),
 ("nearest_update", r"RT_DEV bool nearest_update\("),
])
kern=ranges(os.path.join(ROOT,'gpu_ray_tracing_b200','csrc','rt_kernels.cu'),[
 ("misc (Lanes helpers)", r"^namespace rtk"),
 ("running_mean", r"__device__ __forceinline__ f3 running_mean"),
 ("begin_tile..flush (variant 2 helpers)", r"__device__ __forceinline__ void begin_tile"),
 ("quantise", r"__device__ __forceinline__ unsigned quant1"),
 ("retrace_sequential", r"retrace_sequential\(const GpuSphere"),
 ("prim_exact (loop around the quadratic)", r"__device__ __forceinline__ void prim_exact"),
 ("prim_setup (loop around get_ray)", r"__device__ __forceinline__ void prim_setup"),
 ("prim_tile_bound", r"__device__ __forceinline__ void prim_tile_bound"),
 ("prim_sort_list", r"__device__ __forceinline__ void prim_sort_list"),
 ("prim_scan_list (sign tests)", r"__device__ __forceinline__ void prim_scan_list"),
 ("prim_tree_walk", r"__device__ __forceinline__ void prim_tree_walk"),
 ("finalize/push/resolve", r"__device__ __forceinline__ void finalize_texel"),
 ("prim_finish (epilogue loop, texel store)", r"__device__ __forceinline__ void prim_finish"),
 ("prim_flush", r"__device__ __forceinline__ void prim_flush"),
 ("kernel main loop (tile ticket)", r"primary_hits_kernel\(const __grid_constant__"),
 ("after primary kernel", r"^// Wavefront bounces"),
])
def group(fname,line):
    tab={'rt_core.cuh':core,'rt_kernels.cu':kern}.get(fname)
    if not tab: return fname
    g=None
    for start,name in tab:
        if line>=start: g=name
    return fname+": "+str(g)
agg={}
for f,ln,i,s in rows:
    k=group(f,ln); a=agg.setdefault(k,[0,0]); a[0]+=i; a[1]+=s
tot=sum(v[0] for v in agg.values())
print("covered inst %.1f%%"%tot)
for k,v in sorted(agg.items(), key=lambda kv:-kv[1][0]):
    print(f"{v[0]:6.2f}% inst {v[1]:6.2f}% samples  {k}")
