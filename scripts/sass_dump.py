"""Write the SASS of selected kernels of libveles_b200.so to profiles/sass/<kernel>.sass (instruction text only):
python scripts/sass_dump.py   (no GPU needed)"""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "veles_b200", "lib", "libveles_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
# (file stem, regex on the demangled name) - one representative instantiation per kernel family
WANT = [
    ("gemm_bf16_pair_bn256_kmajor", r"gemm_bf16_tcgen05_kernel<256, false, false, true, false>"),
    ("gemm_bf16_single_bn128_mn_mn", r"gemm_bf16_tcgen05_kernel<128, true, true, false, false>"),
    ("gemm_tf32_single_bn128_kmajor", r"gemm_bf16_tcgen05_kernel<128, false, false, false, true>"),
    ("conv_halo_bn128_forward", r"conv_halo_kernel<128, false>"),
    ("fused_allreduce_gd_multimem", r"fused_allreduce_gd_kernel<true>"),
    ("fused_allreduce_gd_p2p", r"fused_allreduce_gd_kernel<false>"),
    ("gd_update", r"gd_update_kernel"),
    ("gather_rows_u8_bf16", r"gather_rows_kernel<unsigned char, __nv_bfloat16>"),
    ("softmax_eval_f32_bf16", r"softmax_eval_kernel<float, __nv_bfloat16>"),
    ("mse_eval_bf16", r"mse_eval_kernel<__nv_bfloat16>"),
    ("dropout_bf16", r"dropout_kernel<__nv_bfloat16>"),
    ("pool_forward_vec8", r"pool_forward_vec8_kernel"),
    ("pool_backward_vec8", r"pool_backward_vec8_kernel"),
    ("lrn_forward_vec8_n5", r"lrn_forward_vec8_kernel<2>"),
    ("lrn_backward_vec8_n5", r"lrn_backward_vec8_kernel<2>"),
    ("col_sum_partial_vec8", r"col_sum_partial_vec8_kernel"),
    ("space_to_depth_bf16", r"space_to_depth_kernel<__nv_bfloat16>"),
    ("im2col_rowtile", r"im2col_rowtile_kernel"),
    ("xorshift1024star", r"xorshift1024star_kernel"),
    ("pool_fwd_3x3s2_max", r"pool_fwd_3x3s2_kernel<0>"),
    ("pool_bwd_3x3s2", r"pool_bwd_3x3s2_kernel"),
    ("lrn_forward_shfl_n5", r"lrn_forward_shfl_kernel<2, true>"),
    ("lrn_backward_shfl_n5", r"lrn_backward_shfl_kernel<2, true>"),
    ("space_to_depth_rows", r"space_to_depth_rows_kernel"),
    ("col_sum_final", r"col_sum_final_kernel"),
    ("lstm_cell_forward", r"lstm_cell_forward_kernel"),
    ("lstm_cell_backward", r"lstm_cell_backward_kernel"),
    ("sig_conv", r"sig_conv_kernel"),
    ("sig_dwt", r"sig_dwt_kernel"),
    ("sig_peaks_write", r"sig_peaks_write_kernel"),
]


def main():
    os.makedirs(OUT, exist_ok=True)
    text = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, check=True).stdout.decode(errors="replace")
    blocks = re.split(r"\n\s*Function : ", text)
    names = [b.split("\n", 1)[0].strip() for b in blocks[1:]]
    dem = subprocess.run(["c++filt"], input="\n".join(names).encode(), stdout=subprocess.PIPE).stdout.decode().splitlines()
    for stem, pat in WANT:
        rx = re.compile(re.escape(pat).replace("\\ ", " "))
        hit = [i for i, d in enumerate(dem) if rx.search(d)]
        if not hit:
            print("not found:", pat)
            continue
        body = blocks[1 + hit[0]]
        lines = ["// %s" % dem[hit[0]], "// cuobjdump -sass veles_b200/lib/libveles_b200.so (sm_100a), encodings stripped"]
        for ln in body.split("\n"):
            m = re.match(r"\s*/\*([0-9a-f]{4,6})\*/\s+(.*?);\s*/\*", ln)
            if m:
                lines.append("/*%s*/ %s ;" % (m.group(1), re.sub(r"\s+", " ", m.group(2))))
        with open(os.path.join(OUT, stem + ".sass"), "w") as f:
            f.write("\n".join(lines) + "\n")
        print("%-34s %5d instructions" % (stem, len(lines) - 2))


if __name__ == "__main__":
    main()
