"""Build named A/B variants of libb200rt next to the default build (run HERE, nvcc cross-compiles; the .so files travel with gpurun):
    python benchmarks/build_variants.py            # every candidate below
    python benchmarks/build_variants.py stop n48   # some of them
then on the GPU box:  bash benchmarks/ab_lib.sh "" _stop _n48 _n48s _rcp _boxpf _tos     (step time + frame/RNG hash per variant)
Remove them afterwards (cuda-raytracing_b200/lib/libb200rt_*.so): they are git-ignored but ride along in every gpurun snapshot."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _loader import load_build  # noqa: E402

# round-2 candidates, all compile-checked, none measured yet
CANDIDATES = {
    "stop": ["RT_TW_STOPREF=1"],   # end-of-ray marker = valid leaf-flagged reference: no sentinel compare on node iterations
    "n48": ["RT_TW_NODE48=1"],     # 48-byte wide node, bf16 half extents, packed FFMA2 (needs register-pair moves)
    "n48s": ["RT_TW_NODE48=2"],    # same node, scalar FFMAs for the half-extent products
    "rcp": ["RT_TW_RCP_APPROX=1"],  # reciprocal direction (conservative node test only) by MUFU.RCP instead of three IEEE divisions
    "boxpf": ["RT_TW_BOX_PREFILTER=1"],  # exact reference-leaf box check behind a clear-pass filter in approximate arithmetic
    "tos": ["RT_TW_TOS=1", "RT_TW_SPPTR=0"],  # top of the traversal stack in a register (neutral at r01c's balance)
}

if __name__ == "__main__":
    names = sys.argv[1:] or list(CANDIDATES)
    build = load_build()
    for n in names:
        print(n, build.build(defines=CANDIDATES[n], out=f"libb200rt_{n}.so"))
