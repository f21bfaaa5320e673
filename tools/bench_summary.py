"""One-line summaries of bench.py JSON lines (files given on the command line)."""
import json, signal, sys
signal.signal(signal.SIGPIPE, signal.SIG_DFL)   # | head
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable:", e); continue
    if d.get("impl") == "reference":
        print(f, "reference", round(d["value"], 2), d["unit"]); continue
    r = d["roofline"]; al = r.get("all_levels", {})
    print(f, "n_gpus", d["n_gpus"], "value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "pipelined", round(d["e2e_pipelined"]["value"], 1),
          "enc", round(d["encode_mpix_s"], 1), "dec", round(d["decode_mpix_s"], 1),
          "noise", round(d.get("value_noise", {}).get("value", 0), 1), "pageable", {k: round(v, 2) for k, v in d.get("e2e_pageable", {}).items() if isinstance(v, float)},
          "dwt_l1", round(r["frac"], 3), "dwt_all_fwd", round(al.get("forward", {}).get("frac", 0), 3), "dwt_all_inv", round(al.get("inverse", {}).get("frac", 0), 3),
          "clocks", d.get("clocks"))
    print("   alone enc", {k: round(v, 3) for k, v in d["stages_ms_per_frame_alone"].items()})
    print("   alone dec", {k: round(v, 3) for k, v in d["decode_stages_ms_per_frame_alone"].items()})
    print("   batch enc", {k: round(v, 2) for k, v in d["stages_ms_per_frame_in_batch"].items()})
    print("   batch dec", {k: round(v, 2) for k, v in d["decode_stages_ms_per_frame_in_batch"].items()})
