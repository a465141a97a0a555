"""One-line summaries of bench.py JSON lines (stdin or files)."""
import json, sys
for src in (sys.argv[1:] or ["-"]):
    text = sys.stdin.read() if src == "-" else open(src).read()
    for line in text.strip().splitlines():
        if not line.startswith("{"):
            continue
        d = json.loads(line)
        rx, re_ = d.get("roofline_transform") or {}, d.get("roofline_entropy") or {}
        print("%s: %.0f img/s (%.1f %s) e2e %.0f img/s | xf alone %.2f us in-step %s eff %s frac %.3f | pass alone %.2f us in-step %s eff %s | conc %s | launches %s" % (
            src, d.get("images_per_s", 0), d["value"], d["unit"], d["e2e"].get("images_per_s", 0),
            1e3 * rx.get("ms_per_launch", 0), rx.get("ms_per_launch_in_step") and round(1e3 * rx["ms_per_launch_in_step"], 2),
            rx.get("ms_per_launch_effective") and round(1e3 * rx["ms_per_launch_effective"], 2), rx.get("frac", 0),
            1e3 * re_.get("ms_per_launch", 0), re_.get("ms_per_launch_in_step") and round(1e3 * re_["ms_per_launch_in_step"], 2),
            re_.get("ms_per_launch_effective") and round(1e3 * re_["ms_per_launch_effective"], 2),
            d.get("in_step", {}).get("concurrency") and round(d["in_step"]["concurrency"], 2), d.get("gpu_launches")))
        if d.get("e2e_pixels"):
            print("   e2e input: %s" % d["e2e"].get("input", "")[:110])
            print("   e2e_pixels: %.0f img/s | e2e_file_error: %s | e2e parity: %s" % (
                d["e2e_pixels"]["images_per_s"], d.get("e2e_file_error"), str(d["e2e"].get("parity_check"))[:150]))
        for k, v in (d.get("other_configs") or {}).items():
            if "error" in v:
                print("   %s ERROR %s" % (k, v["error"][:300]))
            else:
                r = v.get("roofline") or {}
                print("   %s: value %.1f %s | e2e %.1f %s | roof %s %s | %s | %.1fs" % (
                    k, v["value"], v["unit"], v["e2e"]["value"], v["e2e"]["unit"], r.get("kernel"),
                    r.get("frac") and round(r["frac"], 3), v.get("parity_check", "")[:60], v.get("leg_wall_s", 0)))
                if "e2e_file" in v:
                    print("      e2e_file:", {a: (round(b, 1) if isinstance(b, float) else b) for a, b in v["e2e_file"].items()})
                if "png_file" in v:
                    print("      png_file:", {a: b for a, b in v["png_file"].items() if a != "note"})
        if d.get("one_process"):
            print("   one_process:", d["one_process"])
        if d.get("fast_cpu"):
            print("   fast_cpu:", {k: (round(v["value"], 1), v["unit"]) for k, v in d["fast_cpu"].items() if isinstance(v, dict)})
        if d.get("cpu_baseline"):
            print("   cpu_baseline: %.1f img/s on %d cores" % (d["cpu_baseline"]["images_per_s"], d["cpu_baseline"]["cores"]))
