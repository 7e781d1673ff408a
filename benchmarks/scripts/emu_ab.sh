# NOTE: the B200ATTN_EMU / B200ATTN_SPLIT switches exist only with profiles/r1_s28_prefill_experiments.patch applied
# (git apply profiles/r1_s28_prefill_experiments.patch); the product kernel kept only the variant that won.
# A/B of the FMA-pipe exp2 share in the persistent prefill kernel (B200ATTN_EMU = 0, 1, 2 of every 4 pairs),
# parity of each variant, and an ncu full capture of the persistent kernel for the tensor-pipe figure.
mkdir -p gpurun_out
for e in 0 1 2; do
  B200ATTN_EMU=$e timeout 300 python benchmarks/sweep.py prefill --no-flashinfer > gpurun_out/s28_prefill_emu$e.jsonl 2> gpurun_out/s28_prefill_emu$e.err
  echo "emu=$e rc=$?"; grep -o '"case": "[^"]*", "Hq": [0-9]*\|"frac_measured_bf16": [0-9.]*' gpurun_out/s28_prefill_emu$e.jsonl | paste - - | head -12
done
for e in 1 2; do
  B200ATTN_EMU=$e timeout 600 python -m pytest tests/test_gpu_attention.py -x -q -m gpu -k "prefill" 2>&1 | tail -3
done
for e in 0 2; do
  B200ATTN_EMU=$e timeout 300 ncu --set full --clock-control none --import-source on -k regex:persist -c 2 \
    -o gpurun_out/s28_prefill_persist_emu$e -f python benchmarks/prof_one.py prefill 3 > gpurun_out/s28_ncu_emu$e.log 2>&1
  echo "ncu emu=$e rc=$?"
done
