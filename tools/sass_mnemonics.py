"""Counts of the SASS mnemonics that prove which hardware path a kernel uses (cuobjdump -sass of the engine library):
    python tools/sass_mnemonics.py > profiles/r02_sass_mnemonics.txt"""
import collections
import os
import re
import subprocess

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(R, "aws-k8s-ansible-provisioner_b200", "libb2engine.so")
PATS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "HMMA", "LDGSTS", "SYNCS", "UCGABAR", "ACQBULK", "PREEXIT"]

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
cur, cnt = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        cnt[cur] = collections.Counter()
    elif cur:
        for p in PATS:
            if re.search(r"\b" + p, line):
                cnt[cur][p] += 1
names = list(cnt)
dem = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
print("# SASS mnemonic counts per kernel of libb2engine.so (sm_100a), from: cuobjdump -sass libb2engine.so  (static instruction counts, loops not unrolled)")
print("# UTCHMMA = tcgen05.mma (bf16), UTCBAR = tcgen05.commit, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = TMA tensor load / store (cp.async.bulk.tensor),")
print("# UBLKCP = cp.async.bulk, HMMA = mma.sync, LDGSTS = cp.async, SYNCS = mbarrier ops, UCGABAR = barrier.cluster,")
print("# ACQBULK = griddepcontrol.wait, PREEXIT = griddepcontrol.launch_dependents")
print("# template arguments: linear_tc_kernel<BN tokens, EPI, SMALL, XCP>; linear_pair_kernel<BN, MODE 0 SwiGLU / 1 QKV+RoPE>; linear_resid_kernel<BN>;")
print("# attn_decode_kernel<warps, stages>")
for n, d in zip(names, dem):
    c = cnt[n]
    if sum(c.values()) == 0:
        continue
    d = re.sub(r"^void ", "", d).replace("<unnamed>::", "").replace("(anonymous namespace)::", "")
    if "<" in d and d.index("<") < d.index("("):
        depth = 0
        for i, ch in enumerate(d):
            if ch == "<":
                depth += 1
            elif ch == ">":
                depth -= 1
                if depth == 0:
                    d = d[: i + 1]
                    break
    else:
        d = d.split("(")[0]
    d = d.replace("(int)", "").replace("(bool)", "")
    print(d.ljust(52), " ".join("%s=%d" % (k, c[k]) for k in PATS if c[k]))
