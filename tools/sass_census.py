"""SASS census of the implicit-GEMM kernels:  cuobjdump -sass masterthesis_b200/libmt_b200.so | python tools/sass_census.py"""
import sys, re, collections, subprocess
cur = None
cnt = collections.defaultdict(collections.Counter)
ex = {}
pat = re.compile(r'\b(UTCHMMA|LDTM|UTMALDG|UTMASTG|UTMAREDG|UTCBAR|SYNCS|MEMBAR|HMMA)\b[\w.]*')
for l in sys.stdin:
    m = re.search(r'Function : (\S+)', l)
    if m:
        cur = m.group(1)
        continue
    if cur and 'igemm' in cur:
        for mm in pat.finditer(l):
            op = mm.group(0).split('.')[0]
            cnt[cur][op] += 1
            ex.setdefault((cur, op), l.strip()[:112])
print("cuobjdump -sass masterthesis_b200/libmt_b200.so (sm_100a) - Blackwell-native mnemonics per implicit-GEMM kernel")
print("(UTCHMMA = tcgen05.mma kind::f16, LDTM = tcgen05.ld, UTMALDG / UTMASTG / UTMAREDG = TMA load / store / reduce-add,")
print(" UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, MEMBAR = the cluster syncs of prologue / teardown; no HMMA (legacy mma.sync))")
names = subprocess.run(['c++filt'], input='\n'.join(sorted(cnt)), capture_output=True, text=True).stdout.splitlines()
for f, name in zip(sorted(cnt), names):
    print("\n" + name.split('(')[0])
    print("    " + "  ".join(f"{k}:{v}" for k, v in sorted(cnt[f].items())))
    for k in ('UTCHMMA', 'UTMAREDG'):
        if (f, k) in ex:
            print("      e.g. " + ex[(f, k)])
