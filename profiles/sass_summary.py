"""Writes profiles/sass_r2_{gemm_tc,spmm,p2p}.txt: per-kernel SASS mnemonic counts (tcgen05 = UTCHMMA, TMA = UTMALDG,
tensor-memory loads/stores = LDTM/STTM, 128-bit gathers = LDG.E.128 ...) and a short excerpt, from `cuobjdump -sass`
of the shipped libgdr.so.  Runs without a GPU:  python profiles/sass_summary.py"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "gcn-drug-repurposing_b200", "csrc", "libgdr.so")
WANT = {"gemm_tc": ["gemm_tc_kernel", "gemm_tc_tn_kernel", "gram_topk_tc_kernel"], "spmm": ["spmm_warp_row_kernel"],
        "p2p": ["push_rows_kernel", "mc_broadcast", "mc_reduce"]}
MNEMONICS = ["UTCHMMA", "UTMALDG", "LDTM", "STTM", "UTCBAR", "UTCATOMSWS", "SYNCS", "ELECT", "BRA.U.ANY", "R2UR",
             "LDG.E.128", "LDG.E.64", "STG.E.128", "ST.E.128", "LDS.128", "STS.128", "SHFL", "DFMA", "FFMA", "MULTIMEM",
             "RED", "ATOM", "FMNMX", "SEL"]
INST = re.compile(r"/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)")


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    funcs = re.split(r"\n\s*Function : ", out)[1:]
    for tag, names in WANT.items():
        lines = ["# SASS mnemonic counts per kernel, libgdr.so (sm_100a), `cuobjdump -sass` of the shipped library; round 2", ""]
        for f in funcs:
            name = subprocess.run(["c++filt", f.split("\n", 1)[0].strip()], capture_output=True, text=True).stdout.strip()
            if not any(k in name for k in names):
                continue
            body = [ln for ln in f.splitlines() if INST.search(ln)]
            cnt = collections.Counter()
            for ln in body:
                op = INST.search(ln).group(2)
                for k in MNEMONICS:
                    if op.startswith(k):
                        cnt[k] += 1
            lines.append(f"{name[:150]}\n    instructions {len(body)}  " + "  ".join(f"{k}={v}" for k, v in sorted(cnt.items())))
            key = "UTCHMMA" if cnt["UTCHMMA"] else ("LDG.E.128" if cnt["LDG.E.128"] else None)
            if key:
                i = [k for k, ln in enumerate(body) if key in ln][0]
                lines.append("    excerpt:")
                lines += ["      " + re.sub(r"\s*/\* 0x[0-9a-f]+ \*/", "", ln).strip()[:140] for ln in body[max(0, i - 3):i + 6]]
            lines.append("")
        with open(os.path.join(ROOT, "profiles", f"sass_r2_{tag}.txt"), "w") as fh:
            fh.write("\n".join(lines))


if __name__ == "__main__":
    main()
