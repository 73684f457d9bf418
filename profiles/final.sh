#!/bin/bash
# End-of-round measurement on the GPU box (run through gpurun):  bash profiles/final.sh <tag>
# Everything it writes lands in gpurun_out/; profiles/collect.py turns that into the tracked summaries.
tag=${1:-r02_b}
o=gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -3 | tee $o/${tag}_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -6 | tee $o/${tag}_smoke.txt
python bench.py --impl reference > $o/${tag}_bench_reference.json 2> $o/${tag}_bench_reference.err
python bench.py > $o/${tag}_bench_full.json 2> $o/${tag}_bench_full.err || tail -5 $o/${tag}_bench_full.err
tail -c 600 $o/${tag}_bench_full.json
[ -x profiles/micro/pipe_bench ] || nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/micro/pipe_bench profiles/micro/pipe_bench.cu
profiles/micro/pipe_bench > $o/${tag}_pipe_bench.jsonl
# launch list of the bench command, this library's kernels only (the synthetic-data generation is
# thousands of torch launches); cold-cache, serialised: shares, not absolutes
ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:mfcc|lr_" -c 400 --csv --log-file $o/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --e2e-clips 256 --no-cpu-baseline --stream-steps 20 > /dev/null 2> $o/${tag}_launches.err
# one full capture per fused kernel (300 clips = 1 693 500 frames per launch)
ncu --set full --import-source on --clock-control none -k regex:mfcc_int_fused -s 2 -c 1 -f -o $o/${tag}_int python profiles/run_both.py 300 > /dev/null 2>&1
ncu --set full --import-source on --clock-control none -k regex:mfcc_f32_fused -s 2 -c 1 -f -o $o/${tag}_f32 python profiles/run_both.py 300 > /dev/null 2>&1
ncu --set full --import-source on --clock-control none -k regex:lr_logmel -s 3 -c 1 -f -o $o/${tag}_lr python profiles/run_librosa.py 300 > /dev/null 2>&1
python profiles/run_librosa.py 300 > $o/${tag}_librosa_rate.txt 2>/dev/null
python profiles/accuracy.py > $o/${tag}_accuracy.txt 2>/dev/null
python profiles/stage_rates.py 256 > $o/${tag}_stage_rates.txt 2>/dev/null
[ -x profiles/micro/mma_bench ] || nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/micro/mma_bench profiles/micro/mma_bench.cu
profiles/micro/mma_bench > $o/${tag}_mma_bench.jsonl
ls -la $o | tail -12
