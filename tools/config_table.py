#!/usr/bin/env python
"""Markdown table of bench.py JSON lines (one row per file): workload, GPUs, builds/s, ms/build, stage times, roofline
fractions, clocks record — the table of the five BASELINE configurations in DESIGN.md §6.
    python tools/config_table.py profiles/r02_bench_*.json"""
import json
import sys


def last_json_line(path):
    for ln in reversed(open(path).read().splitlines()):
        if ln.startswith("{"):
            return json.loads(ln)
    raise ValueError("no JSON line in " + path)


def main():
    print("| Workload | GPUs | builds/s | ms/build | e2e builds/s | half-transform | exchange | Coulomb | build vs FP64 peak | clocks (MHz, reasons) | file |")
    print("|---|---|---|---|---|---|---|---|---|---|---|")
    for path in sys.argv[1:]:
        d = last_json_line(path)
        st, rf, ck = d.get("stage_ms", {}), d.get("roofline") or {}, d.get("clocks") or {}
        k = rf.get("kernels", {})

        def tf(name):
            x = k.get(name, {})
            return "%.2f ms, %.1f %s (%.0f %%)" % (x.get("ms", 0), x.get("achieved", 0), "TF/s" if x.get("bound") == "tensor" else "GB/s",
                                                   100 * x.get("frac", 0)) if x else "-"
        print("| %s | %d | %.2f | %.3f | %s | %s | %s | %s | %s | %s/%s %s | `%s` |" % (
            d["config"]["workload"].split(":")[0], d["n_gpus"], d["value"], d["ms_per_step"],
            "%.2f" % d["e2e"]["value"] if d.get("e2e") else "-", tf("half_transform"), tf("exchange"), tf("coulomb"),
            "%.0f %%" % (100 * rf["build_frac_of_fp64_peak"]) if rf.get("build_frac_of_fp64_peak") else "-",
            ck.get("sm_mhz"), ck.get("sm_max_mhz"), ",".join(ck.get("reasons", [])) or "none", path))


if __name__ == "__main__":
    main()
