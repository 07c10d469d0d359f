"""Writes the benchmark scene as committed bytes (tests/golden/random1024_seed1234.bin): rt_create_random_spheres(seed
1234, n 1024), the seeded host generator of SURVEY.md 8(d) config 3 (a host function; no GPU needed).  The CPU arms of
bench.py read this file, so that timing the oracle does not map the product library; tests/test_abi.py checks the file
against the generator.  Run: python tests/golden/make_scene_bins.py"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import gpu_ray_tracing_b200 as rt  # noqa: E402

if __name__ == "__main__":
    scene = rt.create_random_spheres(1024, 1234)
    path = os.path.join(HERE, "random1024_seed1234.bin")
    with open(path, "wb") as f:
        f.write(scene.to_bytes())
    print(path, os.path.getsize(path), "bytes")
