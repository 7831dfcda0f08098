import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; data = rows[2:]
ia, isrc, isamp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot_ex = sum(int(r[iex] or 0) for r in data); tot_s = sum(int(r[isamp] or 0) for r in data)
print("total inst", tot_ex, "samples", tot_s)
by = collections.defaultdict(lambda: [0, 0])
for r in data:
    op = r[isrc].split()[0] if not r[isrc].startswith("@") else r[isrc].split()[1]
    op = op.split(".")[0]
    by[op][0] += int(r[iex] or 0); by[op][1] += int(r[isamp] or 0)
for op, (e, s) in sorted(by.items(), key=lambda kv: -kv[1][0])[:28]:
    print(f"{op:14s} inst {e/tot_ex*100:5.1f}%  samples {s/tot_s*100:5.1f}%")
print("--- top sampled instructions")
for r in sorted(data, key=lambda r: -int(r[isamp] or 0))[:40]:
    st = {hdr[i]: int(r[i] or 0) for i in stall_cols if int(r[i] or 0) > 0}
    top = sorted(st.items(), key=lambda kv: -kv[1])[:3]
    print(r[ia][-5:], f"{int(r[isamp])/tot_s*100:5.2f}%", f"ex {int(r[iex] or 0)/tot_ex*100:4.2f}%", r[isrc][:60], top)
