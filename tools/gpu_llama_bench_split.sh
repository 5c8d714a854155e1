#!/bin/bash
# --split-mode row through the plugin on the GPUs of this box: the reference's llama-bench, weights in the plugin's split buffer type
mkdir -p gpurun_out
OUT=gpurun_out/r2_llama_bench_split.txt
GG=/tmp/tinyllama-random-q4_k_m.gguf
python tools/make_gguf.py $GG > $OUT 2>&1
nvidia-smi -L >> $OUT
NT=$(nproc); [ "$NT" -gt 16 ] && NT=16
export GGML_BACKEND_PATH=$PWD/ggml-neon-opt_b200/lib/libggml-q4k-b200.so
for SM in none row; do
  echo "== plugin, -sm $SM -ngl 99 -nkvo 1, $NT threads" >> $OUT
  (cd oracle/_ref && timeout 600 ./llama-bench -m $GG -p 512 -n 128 -t $NT -ngl 99 -nkvo 1 -sm $SM -r 2) >> $OUT 2>&1
  echo "rc=$?" >> $OUT
done
(cd oracle/_ref && timeout 300 ./llama-bench -v -m $GG -p 0 -n 2 -t $NT -ngl 99 -nkvo 1 -sm row -r 1) > gpurun_out/r2_llama_bench_split_verbose.txt 2>&1
grep -E "graph splits|model buffer size|Split" gpurun_out/r2_llama_bench_split_verbose.txt | sort | uniq -c >> $OUT
tail -30 $OUT
