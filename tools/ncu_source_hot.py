"""Top CUDA source lines of one kernel by warp-stall samples and by executed instructions, from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:K > file.csv`.

    python tools/ncu_source_hot.py file.csv [N]
"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file, hdr, idx, out = None, None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].rsplit("/", 1)[-1]
    elif r[0] == "Line No":
        hdr = r
        idx = {}
        for i, h in enumerate(hdr):
            idx.setdefault(h, i)
    elif hdr and len(r) >= len(hdr) - 2 and r[0].isdigit() and r[2] == "-":
        def num(k):
            try:
                return int(r[idx[k]])
            except Exception:
                return 0
        stalls = {h[6:]: num(h) for h in hdr if h.startswith("stall_") and "Not Issued" not in h}
        out.append((cur_file, int(r[0]), r[1].strip(), num("# Samples"), num("Instructions Executed"), stalls))
tot_s = sum(o[3] for o in out) or 1
tot_i = sum(o[4] for o in out) or 1
print(f"total samples {tot_s}, warp instructions {tot_i}")
print("---- by stall samples")
for f, ln, src, s, ins, st in sorted(out, key=lambda o: -o[3])[:top_n]:
    main = max(st, key=st.get) if st else ""
    print(f"{100*s/tot_s:5.1f}% {100*ins/tot_i:5.1f}%i {main:12s} {f}:{ln:<4d} {src[:110]}")
print("---- by executed warp instructions")
for f, ln, src, s, ins, st in sorted(out, key=lambda o: -o[4])[:top_n]:
    print(f"{100*ins/tot_i:5.1f}%i {100*s/tot_s:5.1f}% {f}:{ln:<4d} {src[:110]}")
