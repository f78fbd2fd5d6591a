"""Concurrency of the fused fold kernel's CTAs from a PLSB_CV_CTA_TRACE dump: how many were resident over time, per SM."""
import sys, collections
rows = [l.strip().split(",") for l in open(sys.argv[1]).read().splitlines()[1:]]
rows = [(int(a), int(b), int(c), int(d)) for a, b, c, d in rows if int(d) > 0]
t0 = min(r[2] for r in rows); t1 = max(r[3] for r in rows)
life = sorted(r[3] - r[2] for r in rows)
print(f"{len(rows)} CTAs, kernel span {(t1-t0)/1e3:.1f} us, CTA life median {life[len(life)//2]/1e3:.1f} us (p10 {life[len(life)//10]/1e3:.1f}, p90 {life[9*len(life)//10]/1e3:.1f})")
ev = sorted([(r[2], 1) for r in rows] + [(r[3], -1) for r in rows])
cur = 0; last = t0; hist = collections.Counter()
for t, d in ev:
    hist[cur] += t - last; last = t; cur += d
tot = sum(hist.values())
avg = sum(k * v for k, v in hist.items()) / tot
print(f"resident CTAs, time-weighted mean {avg:.1f}; share of the span with <=148: {sum(v for k, v in hist.items() if k <= 148)/tot:.2f}, 149-300: {sum(v for k, v in hist.items() if 148 < k <= 300)/tot:.2f}, >300: {sum(v for k, v in hist.items() if k > 300)/tot:.2f}")
nsm = len(set(r[1] for r in rows)); print("SMs used:", nsm)
# ten time slices: mean residency
for i in range(10):
    a = t0 + (t1 - t0) * i / 10; b = t0 + (t1 - t0) * (i + 1) / 10
    occ = sum(max(0, min(r[3], b) - max(r[2], a)) for r in rows) / (b - a)
    print(f"  slice {i}: {occ:.0f} CTAs resident")
