#!/bin/bash
# Build librvcdec.so in-tree for sm_100a (nvcc cross-compiles without a GPU).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../librvcdec.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(${EXTRA_NVCC_FLAGS:-} -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v)
mkdir -p "$HERE/build"
pids=()
for f in aux_kernels conv_f32 conv_mma mrf_fused mrf_pair mrf_pair2 rvcdec synth_front att_mma rvcsyn refine_kernels rvcrefine rmvpe_kernels rmvpe_conv_tc rmvpe_conv_tma rvcrmvpe hubert_kernels gemm_tc gemm_tma hubert_att_tc hubert_att_tma rvchubert; do
  "$NVCC" "${FLAGS[@]}" -c "$HERE/$f.cu" -o "$HERE/build/$f.o" > "$HERE/build/$f.log" 2>&1 &
  pids+=($!)
done
rc=0
for p in "${pids[@]}"; do wait "$p" || rc=1; done
if [ $rc -ne 0 ]; then cat "$HERE"/build/*.log; exit 1; fi
"$NVCC" -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT" "$HERE"/build/aux_kernels.o "$HERE"/build/conv_f32.o "$HERE"/build/conv_mma.o "$HERE"/build/mrf_fused.o "$HERE"/build/mrf_pair.o "$HERE"/build/mrf_pair2.o "$HERE"/build/rvcdec.o "$HERE"/build/synth_front.o "$HERE"/build/att_mma.o "$HERE"/build/rvcsyn.o "$HERE"/build/refine_kernels.o "$HERE"/build/rvcrefine.o "$HERE"/build/rmvpe_kernels.o "$HERE"/build/rmvpe_conv_tc.o "$HERE"/build/rmvpe_conv_tma.o "$HERE"/build/rvcrmvpe.o "$HERE"/build/hubert_kernels.o "$HERE"/build/gemm_tc.o "$HERE"/build/gemm_tma.o "$HERE"/build/hubert_att_tc.o "$HERE"/build/hubert_att_tma.o "$HERE"/build/rvchubert.o -cudart static
echo "built $OUT"
