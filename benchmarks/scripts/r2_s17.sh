#!/bin/bash
# e2e leg through the graph-replay API: full-size line at N=1 (no engine / cpu legs) and a host-bound run (ctx 2048: the
# GPU step is ~25 us, so e2e ms_per_step is the host path per step)
mkdir -p gpurun_out
O=gpurun_out
T=${1:-r2s17}
timeout 600 python bench.py --no-engine --no-cpu-baseline --no-prefill --no-gpu-reference > $O/${T}_bench_full.json 2> $O/${T}_bench_full.err
echo "bench full rc=$?"; python - <<PY
import json
d=json.loads([l for l in open('$O/${T}_bench_full.json') if l.startswith('{')][-1])
print('value',d['value'],d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],d['e2e']['self_check_max_abs_err'],d['e2e']['full_resends_in_timed_steps'])
PY
tail -2 $O/${T}_bench_full.err | cut -c1-300
for extra in "" "--no-graph"; do
timeout 600 python bench.py --ctx 2048 --steps 200 --no-engine --no-cpu-baseline --no-prefill --no-gpu-reference $extra > $O/${T}_bench_host.json 2> $O/${T}_bench_host.err
echo "bench host-bound $extra rc=$?"; python - <<PY
import json
d=json.loads([l for l in open('$O/${T}_bench_host.json') if l.startswith('{')][-1])
print('step ms',d['ms_per_step'],'e2e ms/step',d['e2e']['ms_per_step'],d['e2e']['self_check_max_abs_err'])
PY
done
python - <<'PY'
# host-side profile of the e2e step (cProfile over 300 steps at ctx 2048)
import cProfile, pstats, sys, os, torch, numpy as np
sys.path.insert(0, os.getcwd())
import nanovllmx_b200 as pkg
B,L,ps,Hq,Hkv,D=64,2048,16,32,8,128
dev=torch.device('cuda')
pool=pkg.KVCachePool(torch.bfloat16,dev,(B*(L//ps)+B*40)*ps,1,Hkv,D)
be=pkg.AttentionBackend(pool,Hq,Hkv,D,torch.bfloat16,dev,page_size=ps)
attn=pkg.Attention(Hq,D,None,Hkv,0).to(dev); holder=torch.nn.ModuleList([attn])
class S: pass
seqs=[]
for i in range(B):
    s=S(); s.seq_id=f"r{i}"; s.token_ids=[0]*L; s.kv_indices=list(range(i*(L+640), i*(L+640)+L)); s.cached_kv_len=L-1; seqs.append(s)
class Bt: pass
b=Bt(); b.forward_mode=pkg.ForwardMode.DECODE; b.seqs=seqs; b.num_seqs=B
qkv=torch.randn((B,(Hq+2*Hkv)*D),device=dev,dtype=torch.bfloat16); out=torch.empty((B,Hq*D),device=dev,dtype=torch.bfloat16)
g=pkg.CUDAGraph(); be.prepare_for_cuda_graph_capture(g,B,L+700)
md=be.build_metadata_for_cuda_graph_capture(g,B,L+700)
def fwd(): out.copy_(attn(qkv[:,:Hq*D],qkv[:,Hq*D:(Hq+Hkv)*D],qkv[:,(Hq+Hkv)*D:]))
with pkg.attention_kv_cache(holder,md):
    fwd()
    with g.capture(B): fwd()
def step():
    for s in seqs:
        s.cached_kv_len=len(s.kv_indices); s.token_ids.append(0); s.kv_indices.append(s.kv_indices[-1]+1)
    md=be.build_metadata_for_cuda_graph_replay(g,b,B)
    with pkg.attention_kv_cache(holder,md): g.replay(B)
for _ in range(20): step()
torch.cuda.synchronize()
pr=cProfile.Profile(); pr.enable()
for _ in range(300): step()
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats('tottime').print_stats(14)
PY
