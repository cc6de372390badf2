#!/bin/bash
# tools/ncu_capture.sh -- on a B200 (under gpurun): one `ncu --set full` capture per hot kernel, each only
# after the same command line has exited 0 without ncu (B200_PROFILING.md).  Reports land in
# gpurun_out/ncu/<name>.ncu-rep; tools/ncu_summarize.py turns them into profiles/*.txt on the CPU box.
set -u
mkdir -p gpurun_out/ncu
cap() {   # name  kernel-regex  skip  command...     ($ONLY = substring filter on the name)
    local name=$1 rx=$2 skip=$3; shift 3
    if [ -n "${ONLY:-}" ] && [[ "$name" != *$ONLY* ]]; then return; fi
    "$@" > gpurun_out/ncu/$name.plain.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -f \
        -o gpurun_out/ncu/$name "$@" > gpurun_out/ncu/$name.ncu.log 2>&1
    echo "$name rc=$?"
}
P="python tools/perf.py"
cap dgemm_nn_8192   dgemm_dmma      2 $P gemm f64 N N 8192 8192 8192 3
cap dgemm_nn_1024   dgemm_dmma      2 $P gemm f64 N N 1024 1024 1024 3
cap sgemm_nn_8192   sgemm_tf32x3    2 $P gemm f32 N N 8192 8192 8192 3
cap spack_nn_8192   pack_split      4 $P gemm f32 N N 8192 8192 8192 3
cap zgemm_hn_4096   zgemm_dmma      2 $P gemm c64 H N 4096 4096 4096 3
cap gemv_n_f64      gemv_n_kernel   2 $P gemv f64 N 32768 32768 3
cap gemv_t_f64      gemv_t_kernel   2 $P gemv f64 T 32768 32768 3
cap dot_f64         dot_partial     2 $P dot f64 268435456 3
cap dgemm_nn_500    dgemm_dmma      2 $P gemm f64 N N 500 500 500 3
cap sgemm_mma_1000  sgemm_mma       2 $P gemm f32 N N 1000 1000 1000 3
cap cgemm_mma_500   cgemm_mma       2 $P gemm c32 N N 500 500 500 3
cap gemv_t_slab     gemv_t_slab     2 $P gemv f64 T 2000 3000 3
cap asum_f64        asum_partial    2 $P asum f64 268435456 3
cap amax_f32        iamax_partial   2 $P amax f32 134217728 3
# round 1, second half: 3M paths and the narrow tcgen05 tile
cap r1b_zsplit_4096    split_planes    2 $P gemm c64 H N 4096 4096 4096 3
cap r1b_zcombine_4096  combine_planes  2 $P gemm c64 H N 4096 4096 4096 3
cap r1b_z3m_dgemm_4096 dgemm_dmma      6 $P gemm c64 H N 4096 4096 4096 3
cap r1b_sgemm_t64_1000 sgemm_tf32x3    2 $P gemm f32 N N 1000 1000 1000 3
cap r1b_cpack3_4096    pack_split3     2 $P gemm c32 H N 4096 4096 4096 3
cap r1b_ccombine3_4096 combine3        2 $P gemm c32 H N 4096 4096 4096 3
cap r1b_c3m_sgemm_4096 sgemm_tf32x3    6 $P gemm c32 H N 4096 4096 4096 3
