# NOTE: the B200ATTN_EMU / B200ATTN_SPLIT switches exist only with profiles/r1_s28_prefill_experiments.patch applied
# (git apply profiles/r1_s28_prefill_experiments.patch); the product kernel kept only the variant that won.
# A/B of the prefill pipeline variants (B200ATTN_SPLIT = MODE 0/1/2) and the FMA-pipe exp2 share (B200ATTN_EMU)
mkdir -p gpurun_out
M=${MODES:-"2:0 2:1 0:0"}
B200ATTN_SPLIT=${TESTMODE:-2} timeout 300 python -m pytest tests/test_gpu_attention.py -x -q -m gpu -k "prefill" 2>&1 | tail -15
for cfg in $M; do
  m=${cfg%%:*}; e=${cfg##*:}
  B200ATTN_SPLIT=$m B200ATTN_EMU=$e timeout 300 python benchmarks/sweep.py prefill --no-flashinfer > gpurun_out/s28_prefill_mode${m}_emu$e.jsonl 2> gpurun_out/s28_prefill_mode${m}_emu$e.err
  echo "mode=$m emu=$e rc=$?"; grep -o '"case": "[^"]*", "Hq": [0-9]*\|"frac_measured_bf16": [0-9.]*' gpurun_out/s28_prefill_mode${m}_emu$e.jsonl | paste - - | grep -v "64 seqs" | head -12
  tail -3 gpurun_out/s28_prefill_mode${m}_emu$e.err
done
