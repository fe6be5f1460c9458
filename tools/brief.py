import json, sys
for l in sys.stdin:
    if l.startswith("{"):
        d = json.loads(l); r = d["roofline"]
        print("  ms/step %.3f kernel_ms %.3f evals/s %.3e e2e_ms %.3f pipe_util %.3f chunks %s thr %s PT %s" % (
            d["ms_per_step"], r["kernel_ms"], d["value"], d["e2e"]["ms_per_step"], r["pipe_util"] or 0,
            d["config"]["segment_chunks"], d["config"]["threads_per_cta"], d["config"]["ploidy_tile"]))
    elif "rror" in l:
        print(l.rstrip())
