#!/bin/bash
# Builds oracle/_ref/libref_cuda.so from the reference's OWN CUDA kernel: the CUDA C text that
# src/cuda/mod.rs embeds between `const CUDA_SOURCE: &str = r#"` and `"#;` is extracted where it
# lies under /root/reference (never copied into the tracked tree), our launcher.inc is appended, and
# nvcc compiles the pair for sm_100a.  Runs only where /root/reference exists (the build container);
# the GPU box uses the prebuilt .so that travels with the snapshot (oracle/_ref/ is git-ignored).
set -e
here="$(cd "$(dirname "$0")" && pwd)"
ref=${RRT_REFERENCE:-/root/reference}/src/cuda/mod.rs
out="$here/../_ref"
[ -f "$ref" ] || { echo "ref_cuda: $ref not present, keeping any prebuilt $out/libref_cuda.so"; exit 0; }
mkdir -p "$out"
awk '/const CUDA_SOURCE: &str = r#"/{on=1; next} /^"#;/{on=0} on' "$ref" > "$out/ref_render.cu"
[ -s "$out/ref_render.cu" ] || { echo "ref_cuda: extraction produced nothing" >&2; exit 1; }
{ echo '#include <string.h>'; cat "$out/ref_render.cu"; cat "$here/launcher.inc"; } > "$out/ref_render_with_launcher.cu"
# the reference compiles with NVRTC defaults (no fast-math), src/cuda/mod.rs:345-366
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -Xcompiler -fPIC -shared \
     -o "$out/libref_cuda.so" "$out/ref_render_with_launcher.cu" -cudart shared
rm -f "$out/ref_render_with_launcher.cu"
lines=$(wc -l < "$out/ref_render.cu"); rm -f "$out/ref_render.cu"
echo "built $out/libref_cuda.so ($lines lines of reference CUDA C)"
