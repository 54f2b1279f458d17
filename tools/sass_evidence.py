"""Count the Blackwell-path SASS mnemonics per kernel of libfoc_b200.so and print a few lines of context for each tensor kernel.
  python tools/sass_evidence.py > profiles/r2_sass.md
(tcgen05.mma -> UTCHMMA, tcgen05.ld/st -> LDTM/STTM, cp.async.bulk.tensor -> UTMALDG/UTMASTG, cp.async.bulk -> UBLKCP; HMMA would be the
legacy mma.sync path -- B200_PROFILING.md "What proves a Blackwell-native kernel")."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "focnerf_b200", "libfoc_b200.so")
PAT = re.compile(r"\b(UTCHMMA|UTCQMMA|UTCBAR|UTCCP|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|UTMAPF|HMMA|REDG?|ATOMG?|LDGSTS|SYNCS)\b")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    cur, cnt, ctx, total = None, collections.OrderedDict(), {}, {}
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1); cnt[cur] = collections.Counter(); ctx[cur] = []; total[cur] = 0
            continue
        if cur and re.search(r"/\*[0-9a-f]{4}\*/", line):
            total[cur] += 1
            for mm in PAT.findall(line):
                cnt[cur][mm] += 1
                if mm in ("UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "UBLKCP") and sum(1 for c in ctx[cur] if mm in c) < 2:
                    ctx[cur].append(re.sub(r"\s+", " ", re.sub(r"/\*[0-9a-f]+\*/", "", line)).strip())
    dm = demangle(list(cnt))
    short = lambda n: re.sub(r"\(.*", "", dm.get(n, n).replace("(anonymous namespace)::", "").replace("void ", ""))
    print("# SASS mnemonics per kernel of focnerf_b200/libfoc_b200.so (`cuobjdump -sass`, sm_100a)\n")
    print("`UTCHMMA` = tcgen05.mma (fp16 kind), `LDTM`/`STTM` = tcgen05.ld/st (TMEM), `UTMALDG`/`UTMASTG` = TMA tensor load/store, `UBLKCP` = "
          "cp.async.bulk, `SYNCS` = mbarrier ops, `REDG` = red.global (the hash grid's fp16x2 reductions).  No kernel contains `HMMA` (mma.sync).\n")
    print("| kernel | SASS instructions | UTCHMMA | UTCBAR | LDTM | STTM | UTMALDG | UTMASTG | UBLKCP | SYNCS | REDG | HMMA |\n|" + "---|" * 12)
    keys = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "REDG", "HMMA"]
    for n, c in cnt.items():
        tensorish = any(c[k] for k in keys[:7])
        if tensorish or (c["REDG"] and "__half, 3, 2" in short(n)):      # the reduction kernels: only the instantiation the train step uses
            print(f"| `{short(n)}` | {total[n]} | " + " | ".join(str(c[k]) for k in keys) + " |")
    print("\n## first occurrences in the tensor kernels\n")
    for n, lines in ctx.items():
        if cnt[n]["UTCHMMA"]:
            print(f"`{short(n)}`\n```")
            for l in lines:
                print(l)
            print("```")
    assert not any(c["HMMA"] for c in cnt.values())


if __name__ == "__main__":
    main()
