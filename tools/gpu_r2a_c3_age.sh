#!/bin/bash
# Round 2, first GPU call: the config-3 open item.  ~8 GPU-minutes.  Output: gpurun_out/r2a_c3_age.txt
# Build the shared-memory staging variant HERE first (nvcc cross-compiles; the .so travels with the snapshot):
#   I="-DSRT_INLINE_MATH"; tools/build_variant.sh stage "$I" "$I -DSRT_INLINE_STIX -DSRT_INLINE_PLASMA -DSRT_INLINE_ACCEPT -DSRT_GRID_STAGE"
#   gpurun --timeout 1200 -- 'bash tools/gpu_r2a_c3_age.sh'
set -u
mkdir -p gpurun_out
out=gpurun_out/r2a_c3_age.txt
: > $out
run() { echo "## $*" >> $out; timeout 300 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
# knob parity first (bit-identical results required before any timing counts)
run SRT_X=0 python -m pytest tests/test_gpu_zz_phase_align.py -x -q
# rate against age, default knobs
for ms in 200 1000 4000; do run SRT_X=0 python tools/c3_age.py --maxsteps $ms --save /tmp/c3_$ms.npy; done
# young rays with an old population's cell footprint
run SRT_X=0 python tools/c3_age.py --maxsteps 200 --alts 1000,4000,9000,15000,21000
# ingredients apart
run SRT_X=0 python tools/c3_age.py --maxsteps 1000 --model ngo --igrf 1
run SRT_X=0 python tools/c3_age.py --maxsteps 1000 --model grid --igrf 0
# scheduling knobs
run SRT_TEAM_PHASE_ALIGN=1 python tools/c3_age.py --maxsteps 1000 --compare /tmp/c3_1000.npy
run SRT_TEAM_BLOCK=512 python tools/c3_age.py --maxsteps 1000 --compare /tmp/c3_1000.npy
run SRT_TEAM_PHASE_ALIGN=1 SRT_TEAM_BLOCK=512 python tools/c3_age.py --maxsteps 1000 --compare /tmp/c3_1000.npy
run SRT_KERNEL=thread python tools/c3_age.py --maxsteps 1000 --compare /tmp/c3_1000.npy
# team-cooperative fetch of the cell record through shared memory (build variant, see the header)
if [ -f tools/_variants/stage.so ]; then
  run SRT_X=0 python tools/c3_age.py --maxsteps 200 --lib tools/_variants/stage.so --compare /tmp/c3_200.npy
  run SRT_X=0 python tools/c3_age.py --maxsteps 1000 --lib tools/_variants/stage.so --compare /tmp/c3_1000.npy
  run SRT_TEAM_PHASE_ALIGN=1 python tools/c3_age.py --maxsteps 4000 --lib tools/_variants/stage.so --compare /tmp/c3_4000.npy
fi
# NaN fast-forward (build variant): the whole parity suite against it, then config 3 and config 2 timings
#   tools/build_variant.sh nanff "$I" "$I -DSRT_INLINE_STIX -DSRT_INLINE_PLASMA -DSRT_INLINE_ACCEPT -DSRT_NAN_FASTFWD"
if [ -f tools/_variants/nanff.so ]; then
  run SRT_LIB=$PWD/tools/_variants/nanff.so python -m pytest tests/test_gpu_parity.py tests/test_gpu_grid.py -x -q
  run SRT_X=0 python tools/c3_age.py --maxsteps 4000 --lib tools/_variants/nanff.so --compare /tmp/c3_4000.npy
  run SRT_VARIANT_ONLY=nanff python tools/gpu_variants.py 100 100 20000 team
  run SRT_VARIANT_ONLY=libsrt python tools/gpu_variants.py 100 100 20000 team
fi
tail -60 $out
