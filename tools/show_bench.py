#!/usr/bin/env python3
"""Print the key numbers of bench.py JSON lines:  python tools/show_bench.py gpurun_out/*.json"""
import json
import sys

for path in sys.argv[1:]:
    try:
        d = json.loads(open(path).read().strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        print(path, "unreadable:", e)
        continue
    c, r, e2e = d.get("config", {}), d.get("roofline") or {}, d.get("e2e") or {}
    ph = c.get("phases_ms", {})
    it = max(1, c.get("cg_iterations", 1))
    print("%s: N=%s value %.3f MDOF/s  step %.1f ms  cg %.1f ms (%d it, %.4f ms/it)  elems %.2f  struct %.2f  host asm/solve %.1f/%.1f  "
          "spmv %.4f ms (req %.0f GB/s, frac %.3f)  e2e %.3f MDOF/s  clocks %s\n    %s"
          % (path, d.get("n_gpus"), d.get("value", 0) / 1e6, d.get("ms_per_step", 0), ph.get("cg", 0), it, ph.get("cg", 0) / it,
             ph.get("elements", 0), ph.get("structure", 0), ph.get("host_wall_assemble_call", 0), ph.get("host_wall_solve_call", 0),
             r.get("ms_per_launch", 0), r.get("kernel_requested_gbs", 0), r.get("frac", 0), (e2e.get("value") or 0) / 1e6,
             d.get("clocks"), c.get("parallelism")))
