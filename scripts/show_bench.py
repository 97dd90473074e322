"""One compact line per bench JSON file given on the command line."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "ERR", e); continue
    r = d.get("roofline", {}); p = d.get("parity", {})
    lg = d.get("rmat22") or d.get("weak22") or {}
    print(f.split("/")[-1], "|", d["config"]["workload"].split(":")[0], "| N", d["n_gpus"], "| GTEPS", round(d["value"], 1),
          "(min", round(d.get("gteps_min", 0), 1), "max", round(d.get("gteps_max", 0), 1), ") ms", round(d["ms_per_step"], 3),
          "| frac", round(r.get("frac", 0), 3), "| e2e", round(d["e2e"]["value"], 1), round(d["e2e"].get("ms_per_step") or 0, 2), "ms",
          "| strict", round(d.get("strict_graph500_rule", {}).get("value", 0), 1), "| build_s", round(d.get("build_s", 0), 2),
          "| parity", p.get("roots_checked"), p.get("result"), "bfsCPU roots", p.get("bfs_cpu_roots"),
          "| cpu", (d.get("cpu_baseline") or {}).get("value"), "| legacy", lg.get("workload"), round(lg.get("value", 0), 1), round(lg.get("frac", 0), 3))
