python bench.py --steps 5 --warmup 3 2>/dev/null > gpurun_out/b1.json; python -c "
import json
d=json.loads([l for l in open('gpurun_out/b1.json') if l.startswith('{')][0]); print('1gpu', d['ms_per_step'], d['phases'])"
for k9 in tail overlap; do
DLTI_K9=$k9 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 2>/dev/null > gpurun_out/b2_$k9.json; python -c "
import json
d=json.loads([l for l in open('gpurun_out/b2_$k9.json') if l.startswith('{')][0]); print('2gpu $k9', d['ms_per_step'], d['phases'])"
done
