"""Executed-instruction mix of one kernel launch of an ncu report, by SASS opcode and by the code between block barriers
(= the phases of a tile kernel), plus the per-lane issue loops of bulk copies (executions of UBLKCP / UBLKPF per issue
of the loop head).  usage: python scripts/ncu_sass_mix.py <report.ncu-rep> <kernel regex> [launch index]"""
import collections, csv, re, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "-k", "regex:" + kern,
                      "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = next(r for r in rows if "Instructions Executed" in r)
ie, isrc, ist, ia = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Address")
seen, ins = set(), []
for r in rows:
    if len(r) <= ie or not r[ie].isdigit() or r[ia] in seen:   # (the page lists the function once per cubin copy)
        continue
    seen.add(r[ia])
    s = r[isrc].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", s)
    ins.append(((m.group(2) if m else s).split(".")[0], int(r[ie]), int(r[ist] or 0)))
tot, stt = sum(n for _, n, _ in ins), max(1, sum(s for _, _, s in ins))
print(f"{len(ins)} SASS instructions, {tot} warp-level executions, {stt} stall samples")
by, st = collections.Counter(), collections.Counter()
for op, n, s in ins:
    by[op] += n
    st[op] += s
for op, n in by.most_common(24):
    print(f"  {op:10s} {n / tot * 100:5.1f} % of executed   {st[op] / stt * 100:5.1f} % of stall samples")
seg, acc = [], 0
for op, n, _ in ins:
    acc += n
    if op == "BAR":
        seg.append(acc)
        acc = 0
seg.append(acc)
print("between block barriers (% of executed):", " | ".join(f"{a / tot * 100:.1f}" for a in seg))
