"""SASS excerpts per kernel family out of lib/libdtds_b200.so (cuobjdump -sass; no GPU needed).

    python scripts/sass_excerpts.py        # writes profiles/sass/<family>.txt

For one instance of every kernel family: the instruction histogram and the code around the
instructions that prove the Blackwell path (UTCHMMA/UTCQMMA = tcgen05.mma, UTMALDG = TMA load, LDTM =
tcgen05.ld, UTCBAR = tcgen05.commit, MULTIMEM/LDGMC/REDG.MC... = multimem, SYS-scope LDG/STG = peer
loads and stores).
"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed-training-and-deepspeed_b200", "lib", "libdtds_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")

FAMILIES = [
    # (file name, regex on the demangled kernel name, mnemonics to centre excerpts on)
    ("gemm_1cta_bn256", r"gemm_bf16_kernel<256, false, false, false, 1>", ["UTMALDG", "UTCHMMA", "UTCBAR", "LDTM"]),
    ("gemm_pair_bn256", r"gemm_bf16_kernel<256, false, false, false, 2>", ["UTMALDG", "UTCHMMA", "UTCBAR", "LDTM"]),
    ("gemm_allgather_fused", r"gemm_bf16_kernel<128, false, false, true, 1>", ["LDG.E.128.STRONG.SYS", "UTMALDG", "UTCHMMA"]),
    ("attention_fwd", r"attn_fwd_kernel<4>", ["UTMALDG", "UTCHMMA", "LDTM", "MUFU.EX2"]),
    ("attention_bwd_dkdv", r"attn_bwd_dkdv_kernel<4>", ["UTCHMMA", "LDTM", "STTM"]),
    ("attention_bwd_dq", r"attn_bwd_dq_kernel<4>", ["UTCHMMA", "LDTM"]),
    ("layernorm_fwd", r"ln_fwd_warp_kernel<3>", ["LDG.E.128", "SHFL.BFLY", "MUFU.RSQ"]),
    ("layernorm_bwd_dx", r"ln_bwd_dx_kernel", ["LDG.E.128", "SHFL.BFLY"]),
    ("cross_entropy", r"ce_fwd_bwd_kernel", ["MUFU.EX2", "MUFU.LG2", "STG.E.128"]),
    ("adam_flat", r"adam_flat_kernel<false>", ["LDG.E.128", "MUFU.RSQ", "STG.E.128"]),
    ("embedding_fwd", r"embedding_fwd_kernel", ["LDG.E.128", "STG.E.128"]),
    ("allreduce_twoshot_p2p", r"allreduce_twoshot_kernel<__nv_bfloat16, 2>", ["LDG.E.128.STRONG.SYS", "STG.E.128", "LD.E.STRONG.SYS", "ST.E.STRONG.SYS"]),
    ("allreduce_nvls", r"allreduce_nvls_kernel<true>", ["LDGMC", "MULTIMEM", "REDG", "STG.E.128"]),
    ("reduce_scatter_nvls", r"reduce_scatter_nvls_kernel<true>", ["LDGMC", "MULTIMEM"]),
    ("all_gather_nvls", r"all_gather_nvls_kernel", ["MULTIMEM", "STG", "REDG"]),
    ("zero_rs_adam_ag", r"zero_rs_adam_ag_kernel", ["LDGMC", "MULTIMEM", "LDG.E.128.STRONG.SYS", "MUFU.RSQ"]),
]


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    funcs, name, buf = {}, None, []
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if name:
                funcs[name] = buf
            name, buf = m.group(1), []
        elif name:
            buf.append(line)
    if name:
        funcs[name] = buf
    mangled = list(funcs)
    dem = subprocess.run(["c++filt"], input="\n".join(mangled), capture_output=True, text=True).stdout.splitlines()
    os.makedirs(OUT, exist_ok=True)
    for fname, pat, keys in FAMILIES:
        hit = next((m for m, d in zip(mangled, dem) if re.search(pat, d)), None)
        if hit is None:
            print(f"{fname}: no kernel matches {pat}")
            continue
        d = dem[mangled.index(hit)]
        ins = [l for l in funcs[hit] if re.search(r"/\*[0-9a-f]{4,}\*/", l)]
        ops = [re.sub(r"^\s*/\*[0-9a-f]+\*/\s*(@!?U?P\d+\s+)?", "", l).split(" ")[0].rstrip(";") for l in ins]
        hist = collections.Counter(ops).most_common(24)
        out = [f"kernel: {d[:200]}", f"{len(ins)} SASS instructions", "", "instruction histogram (top 24):"]
        out += [f"  {n:6d}  {op}" for op, n in hist]
        for k in keys:
            idx = next((i for i, l in enumerate(ins) if k in l), None)
            if idx is None:
                continue
            out += ["", f"--- around the first {k} ---"]
            out += [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l) for l in ins[max(0, idx - 6):idx + 10]]
        open(os.path.join(OUT, fname + ".txt"), "w").write("\n".join(out) + "\n")
        print(f"{fname}: {len(ins)} instructions, " + ", ".join(f"{k}={sum(k in l for l in ins)}" for k in keys))


if __name__ == "__main__":
    main()
