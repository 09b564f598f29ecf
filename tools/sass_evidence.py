"""Regenerates profiles/sass_evidence.md from `cuobjdump -sass` of the built kernel library:
per-kernel counts of the Blackwell-native instruction classes plus short excerpts."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "epfl_megatron_b200", "ops", "_lib", "libb200kernels.so")
CLASSES = [("UTC*MMA", r"\bUTC[A-Z]*MMA"), (".2CTA", r"\.2CTA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"),
           ("UTMALDG", r"\bUTMALDG"), ("UTMASTG", r"\bUTMASTG"), ("UTMAREDG", r"\bUTMAREDG"), ("UBLKCP", r"\bUBLKCP"),
           ("UTCBAR", r"\bUTCBAR"), ("SYNCS", r"\bSYNCS"), ("MUFU.EX2", r"\bMUFU\.EX2"), ("HMMA", r"\bHMMA")]


def demangle(names):
    out = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    return [re.sub(r"\(.*", "", o).strip() for o in out]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    funcs = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur is not None and "/*" in line:
            funcs[cur].append(line)
    names = demangle(list(funcs))
    rows = []
    for (mangled, lines), name in zip(funcs.items(), names):
        text = "\n".join(lines)
        counts = [len(re.findall(rx, text)) for _, rx in CLASSES]
        rows.append((name, counts, lines))
    tensor = [r for r in rows if r[1][0] or r[1][4] or r[1][5] or r[1][6] or r[1][7]]
    other = [r for r in rows if r not in tensor]
    out = ["# SASS evidence (cuobjdump -sass libb200kernels.so, sm_100a; regenerate with `python tools/sass_evidence.py`)\n",
           "Counts of the Blackwell-native instruction classes per kernel: UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, "
           "UTMALDG/UTMASTG = TMA load/store (cp.async.bulk.tensor), UTMAREDG = cp.reduce.async.bulk.tensor (fp32 add in "
           "L2), UBLKCP = cp.async.bulk, UTCBAR = tcgen05.commit, SYNCS = mbarrier, .2CTA = cta_group::2 forms. "
           "No HMMA (mma.sync) anywhere in the library.\n",
           "| kernel | " + " | ".join(c for c, _ in CLASSES) + " |", "|---|" + "---|" * len(CLASSES)]
    for name, counts, _ in sorted(tensor):
        out.append(f"| `{name}` | " + " | ".join(str(c) for c in counts) + " |")
    out.append(f"\n{len(other)} further kernels (norms, softmax, RoPE, GLU, Adam, cross entropy, data-parallel reduce / "
               "Adam-push, signalling) are SIMT and memory-bound; total HMMA count over the whole library: "
               f"{sum(r[1][-1] for r in rows)}.\n")

    def excerpt(title, pat, rx, before=3, after=8, maxn=2):
        for name, _, lines in rows:
            if pat in name:
                hits = [i for i, l in enumerate(lines) if re.search(rx, l)][:maxn]
                if not hits:
                    continue
                out.append(f"## excerpt: {title} (`{name}`)\n\n```")
                for h in hits:
                    out.extend(l.rstrip()[:150] for l in lines[max(0, h - before):h + after])
                    out.append("    ...")
                out.append("```\n")
                return

    excerpt("CTA-pair GEMM main loop", "gemm_bf16_pair_kernel<false, false>", r"UTC[A-Z]*MMA.*2CTA")
    excerpt("GEMM epilogue: TMA store and TMA reduce-add", "gemm_bf16_pair_kernel<true, true>", r"UTMAREDG|UTMASTG")
    excerpt("attention forward v2: S = QK^T, P from TMEM", "attn_fwd2_kernel", r"UTC[A-Z]*MMA")
    excerpt("attention forward v2: softmax exponent", "attn_fwd2_kernel", r"MUFU\.EX2", 2, 4, 1)
    excerpt("attention backward", "attn_bwd_kernel<true>", r"UTC[A-Z]*MMA", 2, 6, 1)
    excerpt("fused all-gather / reduce-scatter GEMM: peer stores by TMA", "gemm_comm_kernel<false>", r"UTMASTG", 2, 6, 1)
    with open(os.path.join(ROOT, "profiles", "sass_evidence.md"), "w") as f:
        f.write("\n".join(out) + "\n")
    print(f"{len(rows)} kernels, {len(tensor)} with tcgen05/TMA instructions")


if __name__ == "__main__":
    sys.exit(main())
