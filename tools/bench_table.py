"""Development aid: markdown tables for BASELINE.md from the bench lines under profiles/.

    python tools/bench_table.py profiles/r2_bench_cfg1.json profiles/r2_bench_cfg2.json ...
"""
import json
import sys


def g(x, d=3):
    return ("%." + str(d) + "g") % x


rows, cpu, enc = [], [], []
for path in sys.argv[1:]:
    l = json.loads(open(path).read().strip().splitlines()[-1])
    c = l["config"]
    r, e = l["roofline"], l["e2e"]
    rows.append("| %s | %d | %.2f | %s | %s | %.3f | %.0f | %s | %.3f / %.3f | %.3f | %s | %s | %.1f | %s |" % (
        c.get("config"), l["n_gpus"], c["wire_bytes_per_gpu"] / 1e9, g(c["frames_per_gpu"]), g(l["value"]), l["ms_per_step"],
        l["wire_GBps"], " + ".join("%.2f" % v for v in r["launch_ms"].values()), r["frac"], r["frac_survey_bytes"],
        r["step_frac"] or 0, "%.2f" % r["split_ms_per_step"], g(e["value"]), e["wire_GBps"] / l["n_gpus"],
        ("%.2f" % e["pcie_roofline"]["frac"]) if e.get("pcie_roofline") else "—"))
    cb = l.get("cpu_baseline")
    if cb:
        cpu.append("| %s | %s | %s | %s | %d | %.0f x | %.1f x | %.1f x |" % (
            c.get("config"), g(cb["value"]), g(cb["in_place_value"]), g(cb["single_thread_value"]), cb["cores"],
            l["value"] / cb["value"], cb["e2e_over_faithful"], cb["e2e_over_in_place"]))
    en = l.get("encode")
    if en:
        enc.append("| %s | %s | %.2f | %s | %.0f | %.3f | %s | %s |" % (
            c.get("config"), g(en["packets"]), en["ms"], g(en["packets_per_s"]), en["frames_GBps"], en["roofline"]["frac"],
            g(en["e2e"]["value"]) if "e2e" in en else "—", g(en["cpu_baseline"]["value"]) if "cpu_baseline" in en else "—"))
print("| config | GPUs | GB / GPU | frames / GPU | frames/s (device-resident) | ms/step | wire GB/s | k_decode ms per launch | "
      "k_decode frac of HBM peak (layout / survey bytes) | whole step frac | split ms | frames/s end to end | e2e upload GB/s per GPU | of PCIe peak |")
print("|" + "---|" * 14)
print("\n".join(rows))
if cpu:
    print("\n| config | CPU port faithful, all cores (frames/s) | in place | one thread | cores | device-resident / faithful | e2e / faithful | e2e / in place |")
    print("|" + "---|" * 8)
    print("\n".join(cpu))
if enc:
    print("\n| config | packets | ms (device-resident) | packets/s | frames GB/s | frac of HBM peak | packets/s end to end | CPU encoder, all cores |")
    print("|" + "---|" * 8)
    print("\n".join(enc))
