"""Development aid: the few numbers of a bench.py JSON line one looks at first."""
import json
import sys

for path in sys.argv[1:]:
    try:
        l = json.load(open(path))
    except Exception as e:
        print(path, "unreadable:", e)
        continue
    r, e = l["roofline"], l["e2e"]
    print("%s: %.3f ms/step %.3g %s | k_decode %s frac %.3f (survey bytes %.3f) split %.3f ms | e2e %.1f ms %.1f GB/s | "
          "parity %s mispredicts %s launches %s" % (
              l["config"].get("config"), l["ms_per_step"], l["value"], l["unit"],
              {k: round(v, 3) for k, v in r["launch_ms"].items()}, r["frac"], r["frac_survey_bytes"], r["split_ms_per_step"],
              e["ms_per_step"], e["wire_GBps"], (l.get("parity") or {}).get("ok"), l.get("row_mispredicts"), l.get("gpu_launches")))
