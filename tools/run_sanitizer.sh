#!/bin/bash
# compute-sanitizer over tools/sanitizer_target.py: memcheck, racecheck, initcheck, synccheck.
# Every torch tensor is its own cudaMalloc (no caching allocator), so bounds and initialisation are
# tracked per buffer.  Logs -> gpurun_out/r2_sanitizer_<tool>.txt (copied to profiles/ by hand).
out=${1:-gpurun_out}
mkdir -p "$out"
export PYTORCH_NO_CUDA_MEMORY_CACHING=1
for tool in memcheck racecheck initcheck synccheck; do
    extra=""
    [ "$tool" = memcheck ] && extra="--leak-check no"
    [ "$tool" = initcheck ] && extra="--track-unused-memory no"
    timeout 1500 compute-sanitizer --tool "$tool" $extra --print-limit 20 --log-file "$out/r2_sanitizer_$tool.txt" \
        python tools/sanitizer_target.py > "$out/r2_sanitizer_$tool.stdout" 2>&1
    echo "== $tool: exit $?"; tail -2 "$out/r2_sanitizer_$tool.stdout"; tail -3 "$out/r2_sanitizer_$tool.txt"
done
