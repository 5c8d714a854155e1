#!/bin/bash
# the reference's own headline tool (README.md:189-193) on a random-init TinyLlama-shaped Q4_K_M GGUF:
# on the host cores alone, with the plugin claiming the mul_mats only, and with the plugin claiming the cheap ops around them.
mkdir -p gpurun_out
OUT=gpurun_out/r2_llama_bench.txt
GG=/tmp/tinyllama-random-q4_k_m.gguf
python tools/make_gguf.py $GG > $OUT 2>&1
NT=$(nproc); [ "$NT" -gt 16 ] && NT=16
REPS=${REPS:-3}
echo "== CPU only (no plugin), $NT threads" >> $OUT
(cd oracle/_ref && timeout 600 ./llama-bench -m $GG -p 512 -n 128 -t $NT -r $REPS) >> $OUT 2>&1
echo "rc=$?" >> $OUT
export GGML_BACKEND_PATH=$PWD/ggml-neon-opt_b200/lib/libggml-q4k-b200.so
for CHEAP in 0 1; do
  export Q4K_B200_CHEAP_OPS=$CHEAP
  echo "== plugin loaded (GGML_BACKEND_PATH), Q4K_B200_CHEAP_OPS=$CHEAP, -ngl 99 -nkvo 1, $NT threads" >> $OUT
  (cd oracle/_ref && timeout 600 ./llama-bench -m $GG -p 512 -n 128 -t $NT -ngl 99 -nkvo 1 -sm none -r $REPS) >> $OUT 2>&1
  echo "rc=$?" >> $OUT
  (cd oracle/_ref && timeout 300 ./llama-bench -v -m $GG -p 0 -n 2 -t $NT -ngl 99 -nkvo 1 -sm none -r 1) > gpurun_out/r2_llama_bench_verbose_$CHEAP.txt 2>&1
  echo "verbose rc=$?" >> $OUT
  grep -E "graph splits|graph nodes|model buffer size|compute buffer size" gpurun_out/r2_llama_bench_verbose_$CHEAP.txt | sort | uniq -c >> $OUT
done
tail -45 $OUT
# thread sweep of the plugin run, decode only: the host side of a token is ~45 tiny CPU graph computes, so fewer threads can be faster
echo "== plugin, decode only, host thread sweep" >> $OUT
(cd oracle/_ref && Q4K_B200_CHEAP_OPS=1 timeout 600 ./llama-bench -m $GG -p 0 -n 128 -t 1,2,4,8 -ngl 99 -nkvo 1 -sm none -r 2) 2>&1 | grep -E "tg128|threads" >> $OUT
tail -8 $OUT
