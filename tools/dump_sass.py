#!/usr/bin/env python
"""Writes COMPLETE SASS listings (one file per kernel, every instruction, encodings stripped) of the kernels that carry the
design into profiles/sass/, and refreshes profiles/sass_mnemonics.txt (bench/sass_report.py).

    python tools/dump_sass.py            # needs cuobjdump and c++filt, no GPU
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed_model_parallel_b200", "_lib", "libdmp_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
# file name -> regex on the demangled kernel name (first match wins)
WANT = {
    "gemm_tcgen05_fwd_colstats_bn64_resident_b.sass": r"gemm_tcgen05_kernel<64, false, false, 1, true>",
    "gemm_tcgen05_fwd_colstats_bn128_resident_b.sass": r"gemm_tcgen05_kernel<128, false, false, 1, true>",
    "gemm_tcgen05_dgrad_bn64.sass": r"gemm_tcgen05_kernel<64, false, true, 0, false>",
    "gemm_tcgen05_wgrad_bn128.sass": r"gemm_tcgen05_kernel<128, true, true, 0, false>",
    "conv_igemm_fwd.sass": r"conv_igemm_kernel<",
    "conv_igemm_wgrad.sass": r"conv_wgrad_kernel<",
    "bn_bwd_fused_bf16.sass": r"bn_bwd_fused_kernel<__nv_bfloat16>",
    "bn_bwd_resident.sass": r"bn_bwd_resident_kernel",
    "bn_apply_sums_bf16.sass": r"bn_apply_sums_kernel<__nv_bfloat16>",
    "dw_strip_bf_fwd_stats_s1.sass": r"dw_strip_bf_kernel<__nv_bfloat16, 1, 4, false, true>",
    "dw_wgrad_strip_s1.sass": r"dw_wgrad_strip_kernel<1, 4, false>",
    "stem_wgrad_reg.sass": r"stem_wgrad_reg_kernel<3, 3, 2>",
    "allreduce_twoshot_f32.sass": r"allreduce_twoshot_kernel<float>",
    "allreduce_oneshot_f32.sass": r"allreduce_oneshot_kernel<float>",
    "allreduce_nvls.sass": r"allreduce_nvls_kernel",
    "allreduce_nvls_sgd.sass": r"allreduce_nvls_sgd_kernel",
    "broadcast.sass": r"broadcast_kernel",
    "syncbn_fwd.sass": r"syncbn_fwd_kernel",
    "attention_gemm_batched.sass": r"gemm_tcgen05_kernel<128, false, false, 0, false>",
}


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    funcs, cur, name = {}, None, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(anonymous namespace\)::", "", name)
            cur = funcs.setdefault(name, [])
            continue
        if cur is None:
            continue
        m = re.match(r"^\s*/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\*", line)
        if m:
            cur.append("/*%s*/ %s ;" % (m.group(1), m.group(2).rstrip()))
    os.makedirs(OUT, exist_ok=True)
    for f in os.listdir(OUT):
        if f.endswith(".sass"):
            os.remove(os.path.join(OUT, f))
    for fname, pat in WANT.items():
        hit = next((n for n in funcs if re.search(pat, n)), None)
        if hit is None:
            print("no kernel matches %r" % pat, file=sys.stderr)
            continue
        body = funcs[hit]
        ops = [re.sub(r"^/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?", "", l).split()[0] for l in body]
        key = ["UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "UTCBAR", "SYNCS", "LDGSTS", "LDGMC", "MULTIMEM", "REDG", "ATOMG", "RED"]
        tally = {k: sum(1 for o in ops if o.startswith(k)) for k in key}
        with open(os.path.join(OUT, fname), "w") as fh:
            fh.write("// %s\n// %d instructions (complete listing, encodings stripped); %s\n" % (
                hit, len(body), ", ".join("%s=%d" % (k, v) for k, v in tally.items() if v)))
            fh.write("\n".join(body) + "\n")
        print("%-52s %6d instructions  %s" % (fname, len(body), {k: v for k, v in tally.items() if v}))
    rep = subprocess.run([sys.executable, os.path.join(ROOT, "bench", "sass_report.py")], capture_output=True, text=True).stdout
    open(os.path.join(ROOT, "profiles", "sass_mnemonics.txt"), "w").write(rep)


if __name__ == "__main__":
    main()
