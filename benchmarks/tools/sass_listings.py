"""Write gzip-compressed SASS listings (cuobjdump -sass) of the kernels that carry the training step into profiles/sass/,
plus an index with the Blackwell-specific mnemonics each one contains.

    python benchmarks/tools/sass_listings.py [lib.so] [outdir]
"""
import gzip
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "distributeddataparallel_b200", "lib", "libb200ddp_kernels.so")
out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "profiles", "sass")
os.makedirs(out, exist_ok=True)

# (file stem, regex on the demangled name)
WANT = [
    ("conv_fprop_tc_gemm_kernel_1_0_64", r"tc::gemm_kernel<\(int\)1, \(int\)0, \(int\)64, \(bool\)0>"),
    ("conv_fprop_swapped_tc_gemm_kernel_0_1_128", r"tc::gemm_kernel<\(int\)0, \(int\)1, \(int\)128, \(bool\)0>"),
    ("conv_dgrad_filter3d_tc_gemm_kernel_4_1_128", r"tc::gemm_kernel<\(int\)4, \(int\)1, \(int\)128, \(bool\)0>"),
    ("conv_wgrad_tc_gemm_kernel_3_5_128", r"tc::gemm_kernel<\(int\)3, \(int\)5, \(int\)128, \(bool\)0>"),
    ("gemm2p_persistent_2cta_bf16_256", r"tc::gemm2p_kernel<\(int\)0, \(int\)0, \(int\)256, \(bool\)1"),
    ("gemm2p_persistent_2cta_bf16_wgrad_256", r"tc::gemm2p_kernel<\(int\)3, \(int\)3, \(int\)256, \(bool\)1"),
    ("attention_fwd", r"attn_fwd_kernel"),
    ("attention_bwd", r"attn_bwd_kernel"),
    ("allreduce_nvls_f32", r"allreduce_nvls_kernel<float>"),
    ("allreduce_twoshot_f32_8", r"allreduce_twoshot_kernel<float, \(int\)8"),
    ("allreduce_push_f32", r"allreduce_push_kernel<float>"),
    ("allreduce_wire16_bf16", r"allreduce_wire16_kernel<__nv_bfloat16>"),
    ("broadcast", r"broadcast_kernel"),
    ("bn_fwd_apply", r"bn_fwd_apply_kernel"),
    ("bn_bwd_reduce", r"bn_bwd_reduce_kernel"),
    ("bn_bwd_apply", r"bn_bwd_apply_kernel"),
    ("bn_bwd_cluster", r"bn_bwd_cluster_kernel"),
    ("sgd", r"sgd_kernel"),
    ("conv_ffma_wgrad", r"conv_gemm_kernel<\(int\)2, \(bool\)1, \(bool\)1>"),
]
MNEMONICS = ["UTCHMMA", "UTCQMMA", "UTCMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMAPF", "UTCATOMSWS", "SYNCS", "LDGMC",
             "STGMC", "REDG", "ACQBULK", "UBLKCP", "MEMBAR", "CCTL", "ERRBAR"]

names = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
funcs = re.findall(r"Function : (\S+)", names)
dem = subprocess.run(["cu++filt"] + funcs, capture_output=True, text=True).stdout.splitlines() if funcs else []
index = []
for stem, pat in WANT:
    hit = next((f for f, d in zip(funcs, dem) if re.search(pat, d)), None)
    if hit is None:
        index.append(f"{stem}: not found ({pat})")
        continue
    sass = subprocess.run(["cuobjdump", "-sass", "-fun", hit, lib], capture_output=True, text=True).stdout
    with gzip.open(os.path.join(out, stem + ".sass.gz"), "wt") as fh:
        fh.write(sass)
    ops = re.findall(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", sass, flags=re.M)
    counts = {m: sum(1 for o in ops if o.split(".")[0] == m or o.startswith(m)) for m in MNEMONICS}
    counts = {k: v for k, v in counts.items() if v}
    demangled = dem[funcs.index(hit)]
    index.append(f"{stem}.sass.gz: {len(ops)} instructions; {counts}; {demangled[:140]}")
with open(os.path.join(out, "INDEX.txt"), "w") as fh:
    fh.write("SASS listings (cuobjdump -sass, sm_100a) of the kernels on the training-step path; gunzip to read.\n"
             "Blackwell-specific mnemonics per kernel: UTCHMMA/UTCMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA load, "
             "UTCBAR = tcgen05.commit, LDGMC/STGMC/REDG...MC = multimem (NVLS), SYNCS = mbarrier.\n\n")
    fh.write("\n".join(index) + "\n")
print("\n".join(index))
