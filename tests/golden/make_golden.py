"""Generates tests/golden/*.npz from the CPU oracle (oracle/librt_oracle.so).

The reference holds no golden vectors for this path and cannot be
built here (Rust + rust-gpu), so these fixtures are REGRESSION vectors of the
oracle restatement itself, pinned by the analytic known-answer tests in
tests/test_oracle.py.  Re-run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

import gpu_raytracer_b200 as g  # noqa: E402
import oracle_binding as ob  # noqa: E402

CASES = {
    "default_160x90": (lambda: g.default_scene(), 160, 90, None),
    "default_wavefront_96x54": (lambda: g.default_scene(), 96, 54, (0, 4, 7)),
    "terrain_40x30_chunkedbvh_128x72": (lambda: g.terrain_scene(40, 30, strategy=g.abi.BVH_CHUNKED), 128, 72, None),
    "terrain_24x20_standardbvh_130x70": (lambda: g.terrain_scene(24, 20, strategy=g.abi.BVH_STANDARD), 130, 70, None),
    "cornell_96x96": (lambda: g.cornell_scene(), 96, 96, None),
}


def run(name):
    mk, w, h, wf = CASES[name]
    r = ob.render(mk(), w, h, wavefront=wf, threads=1)
    return dict(hits=r["hits"], composite=r["composite"], radiance_bits=r["radiance"].view(np.uint32),
                ch0=r["channels"][0], ch1=r["channels"][1], ch2=r["channels"][2])


if __name__ == "__main__":
    for name in CASES:
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **run(name))
        print("wrote", name)
