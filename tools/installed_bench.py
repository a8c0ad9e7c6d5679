#!/usr/bin/env python3
"""The reference's own benchmarks/bench_hypervisor.py (the copy under baseline/_ref, unmodified) run twice on this
box: as it is (CPU), and with the GPU classes installed underneath it (agent_hypervisor_b200.install()).  Every one
of its cases is a single small call, i.e. latency-bound: this is the honest cost of "no CPU fallback" on
call-at-a-time use, to be read beside the batched numbers of bench.py.  Writes gpurun_out/installed_bench.json."""
import json, os, runpy, shutil, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = os.path.join(ROOT, "baseline", "_ref")
script = os.path.join(REF, "benchmarks", "bench_hypervisor.py")
results = os.path.join(REF, "benchmarks", "results", "benchmarks.json")
sys.path.insert(0, os.path.join(REF, "src"))


def run_once():
    runpy.run_path(script, run_name="__main__")
    return {r["name"]: {k: r[k] for k in ("mean_us", "median_us", "p99_us", "ops_per_sec")} for r in json.load(open(results))["benchmarks"]}


out = {"reference_cpu": run_once()}
for name in [m for m in sys.modules if m == "hypervisor" or m.startswith("hypervisor.")]:
    del sys.modules[name]                              # a fresh import, so that the script binds the installed classes
import hypervisor  # noqa: E402
from agent_hypervisor_b200 import install  # noqa: E402
out["rebound"] = install()
out["installed_gpu_classes"] = run_once()
out["ratio_installed_over_reference_mean_us"] = {k: out["installed_gpu_classes"][k]["mean_us"] / max(out["reference_cpu"][k]["mean_us"], 1e-9)
                                                 for k in out["reference_cpu"]}
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "installed_bench.json"), "w"), indent=1)
print(json.dumps(out["ratio_installed_over_reference_mean_us"], indent=1))
