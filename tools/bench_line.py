import json, sys
tag = sys.argv[1]
for n in ("4k", "1080p8x8"):
    try:
        d = json.load(open("gpurun_out/bench_%s_%s.json" % (n, tag))); r = d["roofline"]
        print(n, "value %.0f Mpix/s fwd %.3f ms (%.0f GB/s, %.1f%%) inv %.3f ms (%.0f GB/s, %.1f%%)" % (
            d["value"], r["fwd_ms"], r["fwd_GBps"], 100*r["fwd_GBps"]/r["peak"], r["inv_ms"], r["inv_GBps"], 100*r["inv_GBps"]/r["peak"]))
    except Exception as e:
        print(n, e)
