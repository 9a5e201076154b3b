#!/usr/bin/env bash
# Builds gpax_b200/_lib/libgpax_b200.so for sm_100a (cross-compiles without a GPU).
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
out="$here/../_lib"
mkdir -p "$out" "$here/obj"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v)
pids=()
for f in api gram chol mll predict acq mll_grad draw predict_i8; do
  [ -f "$here/$f.cu" ] || continue
  # incremental: skip objects newer than their source, every header and this script
  if [ -z "${GPX_REBUILD:-}" ] && [ -f "$here/obj/$f.o" ] && \
     [ -z "$(find "$here" "$here/../../include" -maxdepth 1 \( -name "$f.cu" -o -name '*.cuh' -o -name '*.h' -o -name build.sh \) -newer "$here/obj/$f.o" | head -1)" ]; then
    continue
  fi
  extra=()
  # acquisition arithmetic must match NumPy's un-fused mul/add bit for bit
  [ "$f" = acq ] && extra=(-fmad=false)
  [ "$f" = predict ] && [ -n "${GPX_PREDICT_DEBUG_VARIANTS:-}" ] && extra=(-DGPX_PREDICT_DEBUG_VARIANTS)
  ( "$NVCC" "${FLAGS[@]}" "${extra[@]}" -c "$here/$f.cu" -o "$here/obj/$f.o" > "$here/obj/$f.log" 2>&1 \
      || { cat "$here/obj/$f.log"; exit 1; } ) &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
"$NVCC" -shared -o "$out/libgpax_b200.so" "$here"/obj/*.o -lcudart
echo "built $out/libgpax_b200.so"

# jax.ffi handlers (xla_ffi_shim.cc): only where JAX -- and with it the XLA FFI headers -- exists.
# JAX is absent from this image and from the GPU box, so this rule is normally skipped.
if ffi_inc="$(python -c 'import jax; print(jax.ffi.include_dir())' 2>/dev/null)" && [ -d "$ffi_inc" ]; then
  "${CXX:-g++}" -O2 -std=c++17 -fPIC -shared -I"$ffi_inc" -I/usr/local/cuda/include \
      "$here/xla_ffi_shim.cc" -o "$out/libgpax_b200_xla.so" \
      -L"$out" -lgpax_b200 -Wl,-rpath,'$ORIGIN' -L/usr/local/cuda/lib64 -lcudart
  echo "built $out/libgpax_b200_xla.so (jax.ffi handlers)"
else
  echo "jax not importable: skipped xla_ffi_shim.cc (compile-gated, see INTEGRATION.md section 3)"
fi
