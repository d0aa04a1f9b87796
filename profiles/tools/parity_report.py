"""Teacher-forced DR-fixture parity of the CUDA library currently at go2_b200/libgo2_b200.so, as a table (variant builds are copied
over it on the GPU box for A/B runs).  usage: python profiles/tools/parity_report.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "go2-sim2real-locomotion-rl_b200"), os.path.join(ROOT, "tests")]
import test_golden_dr as T  # noqa: E402

for name, kind, budget in T.CASES:
    be, bf, worst, n = T._run(name, kind, "cuda", None)
    print(f"{name:16s} env-ticks {n:6d}  exact mismatches {len(be):3d}  outside band {len(bf):4d} (budget {budget})  worst {worst[0]:9.2f} x band at {worst[1]}")
    if be:
        print("   exact:", be[:6])
