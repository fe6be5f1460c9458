import json, sys
for l in sys.stdin:
    if l.startswith("{"):
        d = json.loads(l); r = d["roofline"]; c = d.get("config_detail", {})
        sp = d.get("e2e_single_process")
        print("  N=%d ms/step %.3f kernel_ms %.3f evals/s %.3e e2e_ms %.3f%s pipe_util %.3f frac %.3f chunks %s PT %s assembly %s" % (
            d["n_gpus"], d["ms_per_step"], r["kernel_ms"], d["value"], d["e2e"]["ms_per_step"],
            (" single-process e2e %.3f (identical %s)" % (sp["ms_per_step"], sp["identical_to_per_rank_result"])) if sp else "",
            r["pipe_util"] or 0, r["frac"] or 0, c.get("segment_chunks"), c.get("ploidy_tile"), c.get("assembly")))
    elif "rror" in l:
        print(l.rstrip())
