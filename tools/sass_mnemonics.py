#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-specific / communication SASS mnemonics in the native library
(`cuobjdump -sass`).  Usage: sass_mnemonics.py [lib.so] > profiles/sass/mnemonics.md"""
import collections
import re
import subprocess
import sys

PAT = re.compile(r"\b(UTCHMMA[\w.]*|UTMALDG[\w.]*|UTMASTG[\w.]*|UTCBAR[\w.]*|UTCATOMSWS[\w.]*|LDTM[\w.]*|UCGABAR_\w+|"
                 r"SYNCS[\w.]*|RED\.[\w.]*|ATOM[\w.]*|REDG[\w.]*|LDG\.E[\w.]*\.SYS|STG\.E[\w.]*\.SYS|"
                 r"STG\.E\.ENL2\.256|LDG\.E\.ENL2\.256[\w.]*|MUFU\.TANH|LDGMC[\w.]*|STGMC[\w.]*|REDGMC[\w.]*|STG\.E\.MC[\w.]*)")


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else "commefficient_b200/lib/libceb200.so"
    out = subprocess.run(["cuobjdump", "-sass", lib], stdout=subprocess.PIPE, text=True).stdout
    try:        # readable kernel names
        out = subprocess.run(["c++filt"], input=out, stdout=subprocess.PIPE, text=True).stdout
    except Exception:
        pass
    per = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (.+)$", line)
        if m:
            cur = m.group(1)
            per[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = PAT.search(line)
        if m:
            per[cur][m.group(1).rstrip(".")] += 1
    print("# SASS evidence (cuobjdump -sass %s, sm_100a)\n" % lib)
    print("Counts of the Blackwell-specific / communication mnemonics per kernel.  `UTCHMMA` = tcgen05.mma "
          "(`.2CTA` = cta_group::2), `UTMALDG` = TMA tensor load (`.2CTA` = completion on the pair leader's "
          "barrier), `LDTM` = tcgen05.ld (TMEM), `UTCBAR` = tcgen05.commit (`.2CTA.MULTICAST` = arrive in both "
          "CTAs), `UTCATOMSWS` = tcgen05.alloc/dealloc, `UCGABAR_*` = cluster barrier, `SYNCS` = mbarrier, "
          "`STG.E.ENL2.256` = 256-bit global store, `RED*` = reduction atomics, `*.SYS` = system-scope (peer) "
          "accesses, `LDGMC...ADD` = multimem.ld_reduce (in-switch reduction, NVLS), multimem.st assembles to a plain `STG.E.128.STRONG.SYS` (the "
          "multicast is a property of the address), `MUFU.TANH` = tanh.approx.  `tcgen05.mma.kind::tf32` assembles to the same `UTCHMMA` mnemonic as "
          "`kind::f16` (the operand kind travels in the instruction descriptor): the TF32 kernels are the "
          "instantiations whose last template argument is 1 (`tc_tile_kernel<..., 1>`, `tc_wgrad_kernel<..., 1>`).\n")
    for name, cnt in per.items():
        if not cnt:
            continue
        print("* `%s`\n    %s" % (name, ", ".join("%s x%d" % kv for kv in sorted(cnt.items()))))


if __name__ == "__main__":
    main()
