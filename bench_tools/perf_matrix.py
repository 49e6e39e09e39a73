"""Step time under every kernel-selection switch, one table (reproduces the "measured neutral / slower" claims of
profiles/conv_bound_analysis.md and DESIGN.md on the box at hand).

    python bench_tools/perf_matrix.py [--steps 200] [--only DPS_HALO2=1,DPS_PAIR=0]

Each row is a fresh `bench.py --no-e2e` process (1 GPU, device-timed CUDA-graph replays) with one environment switch:
  DPS_HALO=0     tap-loop kernel instead of the layer-1 halo kernel            DPS_WHALO=0  plain wgrad kernel
  DPS_HALO2=1    layer-2 halo kernel (resident 64-wide weight slices)          DPS_PAIR=0   no CTA-pair kernel
  DPS_CLUSTER=0  no TMA-multicast clusters for the 256/512-wide layers         DPS_PDL=0    no programmatic dependent launch
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SWITCHES = ["", "DPS_HALO=0", "DPS_WHALO=0", "DPS_HALO2=1", "DPS_PAIR=0", "DPS_CLUSTER=0", "DPS_PAIR=0,DPS_CLUSTER=0",
            "DPS_PDL=0"]


def main() -> int:
    p = argparse.ArgumentParser()
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=10)
    p.add_argument("--only", default=None, help="comma-separated list of switch sets separated by ';'")
    a = p.parse_args()
    sets = a.only.split(";") if a.only else SWITCHES
    rows = []
    for sw in sets:
        env = dict(os.environ)
        for kv in [t for t in sw.split(",") if t]:
            k, v = kv.split("=")
            env[k] = v
        r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", str(a.steps), "--warmup",
                            str(a.warmup), "--no-e2e"], env=env, capture_output=True, text=True, cwd=ROOT)
        line = (r.stdout.strip().splitlines() or ["{}"])[-1]
        try:
            j = json.loads(line)
            rows.append((sw or "(default)", j.get("ms_per_step"), j.get("value"), j.get("ps_error_bits")))
        except json.JSONDecodeError:
            rows.append((sw or "(default)", None, None, r.stderr[-200:]))
        print("%-28s %10s ms/step %12s img/s  err=%s" % rows[-1], flush=True)
    base = rows[0][1] if rows and rows[0][1] else None
    if base:
        print("\ndelta vs %s:" % rows[0][0])
        for sw, ms, _, _ in rows[1:]:
            if ms:
                print("  %-28s %+7.1f us/step" % (sw, (ms - base) * 1e3))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
