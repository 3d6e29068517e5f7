#!/usr/bin/env python3
"""Opcode counts per kernel of the shipped library: python tools/sass_summary.py > profiles/rNN_sass_opcode_summary.txt
(cuobjdump -sass; the mnemonics that prove tcgen05 / TMEM / TMA are named in the header line)."""
import collections, os, re, subprocess, sys

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                         "zvec_b200", "lib", "libzvec_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
names = {}
cur, counts, order = None, {}, []
for line in sass.splitlines():
    m = re.match(r"\s+Function : (\S+)", line)
    if m:
        cur = m.group(1); counts[cur] = collections.Counter(); order.append(cur); continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        counts[cur]["_n"] += 1
        for key in ("UTCHMMA", "UTCIMMA", "UTCQMMA", "LDTM", "UTCBAR", "UTMALDG", "UBLKCP", "SYNCS", "IDP.4A", "HMMA", "IMMA",
                    "LDG", "LDS", "STS", "SHFL", "VOTE", "ATOM", "RED", "CCTL", "UTMAPF"):
            if op.startswith(key):
                counts[cur][key] += 1
dem = subprocess.run(["cu++filt"] + order, capture_output=True, text=True).stdout.splitlines() if order else []
print("# cuobjdump -sass %s (sm_100a): opcode counts per kernel." % os.path.relpath(lib))
print("# UTCHMMA / UTCIMMA = tcgen05.mma kind::f16|tf32 / kind::i8, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UTMALDG = "
      "cp.async.bulk.tensor (TMA),")
print("# UBLKCP = cp.async.bulk, SYNCS = mbarrier ops, IDP.4A = dp4a.  HMMA / IMMA (mma.sync) would be listed if present.")
tot = collections.Counter()
for i, fn in enumerate(order):
    name = dem[i] if i < len(dem) else fn
    name = re.sub(r"zvb::|\(anonymous namespace\)::|<unnamed>::|gemv::", "", name)
    name = re.sub(r"\((int|bool)\)", "", name)          # <(int)2, (bool)1> -> <2, 1>
    name = re.sub(r"\([^<>]*\)$", "", name)              # the parameter list
    c = counts[fn]
    tot.update({k: v for k, v in c.items() if k != "_n"})
    print("%-70s %6d instr  %s" % (name[:70], c["_n"], " ".join("%s=%d" % (k, v) for k, v in c.items() if k != "_n")))
print("# total over %d kernels: %s" % (len(order), " ".join("%s=%d" % kv for kv in sorted(tot.items()))))
