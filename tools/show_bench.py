"""Print the interesting parts of a bench JSON line (argument: file)."""
import json, sys
d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
print("value", d["value"], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["value"], "full", d["e2e"].get("value_full_squares"),
      "diff", d["e2e"].get("max_abs_diff_vs_resident_path"), "six", d["value_six_digit_mode"]["value"])
print("three_g", {k: v for k, v in d.get("value_three_g_shared", {}).items() if k != "note"})
print(json.dumps(d["roofline"], indent=1)[:3500])
print(json.dumps(d.get("hp_grid"), indent=1))
print(json.dumps(d.get("c5_universe"), indent=1)); print(json.dumps(d.get("hbm_rows"), indent=1))
print(json.dumps(d["stages"], indent=1))
print(d["details"])
print(d["cpu_baseline"], d["clocks"], d["gpu_launches"])
