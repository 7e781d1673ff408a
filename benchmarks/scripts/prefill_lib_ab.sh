# Same-box A/B of two builds of libb200attn.so on the prefill sweep: benchmarks/_ab/head.so (previous commit)
# against benchmarks/_ab/new.so (working tree).  Both are built here (make in nano-vllm-x_b200/csrc of each tree)
# and travel with the snapshot; box-to-box variance of the sweep is +-4 %, so only same-box pairs are compared.
mkdir -p gpurun_out
L=nano-vllm-x_b200/libb200attn.so
i=0
for v in new head new head; do
  i=$((i+1)); cp benchmarks/_ab/$v.so $L
  timeout 300 python benchmarks/sweep.py prefill --no-flashinfer > gpurun_out/s38b_prefill_${v}_run$i.jsonl 2> gpurun_out/s38b_prefill_${v}_run$i.err
  echo "$v rc=$? $(grep -o '"frac_measured_bf16": [0-9.]*' gpurun_out/s38b_prefill_${v}_run$i.jsonl | cut -c24-28 | paste -sd' ')"
done
cp benchmarks/_ab/new.so $L
