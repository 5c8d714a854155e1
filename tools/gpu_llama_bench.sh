#!/bin/bash
# the reference's own headline tool (README.md:189-193) on a random-init TinyLlama-shaped Q4_K_M GGUF,
# once on the host cores alone and once with the plugin loaded through GGML_BACKEND_PATH.
mkdir -p gpurun_out
OUT=gpurun_out/r2_llama_bench.txt
python tools/make_gguf.py /tmp/tinyllama-random-q4_k_m.gguf > $OUT 2>&1
NT=$(nproc); [ "$NT" -gt 16 ] && NT=16
echo "== CPU only (no plugin), $NT threads" >> $OUT
(cd oracle/_ref && timeout 600 ./llama-bench -m /tmp/tinyllama-random-q4_k_m.gguf -p 512 -n 128 -t $NT -r 3) >> $OUT 2>&1
echo "rc=$?" >> $OUT
export GGML_BACKEND_PATH=$PWD/ggml-neon-opt_b200/lib/libggml-q4k-b200.so
echo "== plugin loaded (GGML_BACKEND_PATH), -ngl 99 -nkvo 1, $NT threads" >> $OUT
(cd oracle/_ref && timeout 600 ./llama-bench -m /tmp/tinyllama-random-q4_k_m.gguf -p 512 -n 128 -t $NT -ngl 99 -nkvo 1 -sm none -r 3) >> $OUT 2>&1
echo "rc=$?" >> $OUT
echo "== split count (verbose run, 2 tokens)" >> $OUT
(cd oracle/_ref && timeout 300 ./llama-bench -v -m /tmp/tinyllama-random-q4_k_m.gguf -p 0 -n 2 -t $NT -ngl 99 -nkvo 1 -sm none -r 1) > gpurun_out/r2_llama_bench_verbose.txt 2>&1
echo "rc=$?" >> $OUT
grep -E "graph splits|graph nodes|offload|Q4K_B200|buffer size" gpurun_out/r2_llama_bench_verbose.txt | head -40 >> $OUT
tail -40 $OUT
