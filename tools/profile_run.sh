#!/bin/bash
# One GPU-box session that produces everything under profiles/ for a round:
#   gpurun --timeout 1500 -- 'bash tools/profile_run.sh r2'
# (tests, the default bench line, the ncu launch list of a short bench run, full ncu captures of the
# benchmarked convert kernel / the scale kernel / a single-session launch, the no-alloc timing table).
# ncu runs come AFTER the same command has exited 0 without ncu; numbers printed under ncu are never bench values.
R=${1:-r2}
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/${R}_pytest.log 2>&1; tail -3 $O/${R}_pytest.log
python bench.py > $O/${R}_bench.log 2> $O/${R}_bench.err; echo "bench rc=$?"; tail -c 600 $O/${R}_bench.err
python tools/noalloc_timing.py > $O/${R}_noalloc_timing.md 2> $O/noalloc.err; cat $O/${R}_noalloc_timing.md
B="python bench.py --steps 20 --warmup 3 --no-configs --no-side --no-cpu-baseline"
$B > $O/${R}_bench_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${R}_launches.csv $B > $O/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sk_convert_kernel -s 72 -c 1 -o $O/${R}_convert_full $B > $O/ncu_f.log 2>&1
ncu --set full --clock-control none -k regex:sk_convert_kernel -s 8 -c 1 -o $O/${R}_convert_single \
    python tools/run_config.py 3840 2160 3840 2160 NV12 1 4 > $O/ncu_s.log 2>&1
# (what comes back from the box is capped at 64 MiB: the single-session report travels as its digest only)
python tools/ncu_digest.py $O/${R}_convert_single.ncu-rep 3 > $O/${R}_convert_single_digest.txt 2>&1 && rm -f $O/${R}_convert_single.ncu-rep
ncu --set full --clock-control none --import-source on -k regex:sk_scale_tiled -s 70 -c 1 -o $O/${R}_scale_full \
    python tools/run_config.py 2560 1440 1920 1080 I420 32 4 > $O/ncu_sc.log 2>&1
ls -la $O/*.ncu-rep | tail -5
# other scale geometries, device-resident (profiles/r2_scale_kernel.md, bottom table)
for c in "2560 1440 1920 1080 NV12 32" "3840 2160 1920 1080 NV12 16" "3840 2160 2560 1440 NV12 16" "1920 1080 1280 720 NV12 32" \
         "1920 1080 3840 2160 NV12 8" "1280 720 1920 1080 I420 16"; do
    echo "$c: $(python tools/run_config.py $c 30)"
done > $O/${R}_scale_geometries.txt 2>&1; cat $O/${R}_scale_geometries.txt
