#!/bin/bash
# End-of-round measurements on one GPU: the default bench line, the reference arm, prompt-evaluation times, the in-kernel layer
# timeline, and (after the plain runs exited 0) the ncu launch list (caches NOT flushed between kernels: every GEMV pulls the head of the
# next kernel's weights into L2, and a flush would make those bytes count twice) and one full capture of the prompt GEMM.
cd "$(dirname "$0")/.."
O=gpurun_out/r02_final
mkdir -p $O
python bench.py --steps 10 --warmup 5 > $O/bench_default.json 2> $O/bench_default.err; echo "bench rc=$?"
python bench.py --impl reference --steps 10 --warmup 3 > $O/bench_reference.json 2> $O/bench_reference.err; echo "reference rc=$?"
for m in llama3.1:8b qwen2:1.5b gemma:7b; do python tools/bench_prefill.py $m 32,64,512,2000; done > $O/prefill_tokens_per_s.txt 2>&1
python tools/prof_pipeline.py llama3.1:8b spread > $O/pipeline_timeline.txt 2>&1
python bench.py --no-cpu-baseline --no-extras --qtype q4_k --head-qtype q6_k --steps 3 --warmup 2 > $O/bench_q4k_q6k.json 2>/dev/null
python bench.py --steps 1 --warmup 1 --new-tokens 16 --no-cpu-baseline --no-extras > $O/plain_decode.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -s 3000 -c 500 --csv --log-file $O/launches_raw.csv \
    python bench.py --steps 1 --warmup 1 --new-tokens 16 --no-cpu-baseline --no-extras > $O/ncu_decode.log 2>&1
python tools/ncu_prefill.py llama3.1:8b 64 > $O/plain_prefill.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_prefill_gemm -s 64 -c 4 -o $O/prefill_gemm_full python tools/ncu_prefill.py llama3.1:8b 64 > $O/ncu_prefill.log 2>&1
ls -la $O
