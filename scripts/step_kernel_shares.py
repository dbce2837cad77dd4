"""Aggregate scripts/step_profile.py output (CUPTI records of one replayed step) by kernel name: python scripts/step_kernel_shares.py file"""
import re, sys, collections
tot = collections.defaultdict(lambda: [0, 0.0])
span0, span1 = None, 0.0
for line in open(sys.argv[1]):
    if line.startswith("#") or not line.strip():
        continue
    p = line.split(None, 3)
    if len(p) < 4:
        continue
    st, du, en, nm = float(p[0]), float(p[1]), float(p[2]), p[3].strip()
    nm = re.sub(r"<.*", "", nm) if not nm.startswith("gemm_kernel") and not nm.startswith("void") else nm
    nm = nm.replace("void ", "")
    nm = re.sub(r"\(.*", "", nm)
    tot[nm][0] += 1; tot[nm][1] += du
    span0 = st if span0 is None else min(span0, st); span1 = max(span1, en)
s = sum(v[1] for v in tot.values())
print(f"step span {span1 - span0:.1f} us, sum of kernel times {s:.1f} us")
for nm, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{t:10.1f} us {100 * t / s:5.1f}%  x{n:<4d} {nm[:100]}")
