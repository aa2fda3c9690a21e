TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
i=0
for V in "ORBX_GATHER_CTAS=4" "ORBX_GATHER_CTAS=12" "ORBX_GATHER_COPY=1" "ORBX_GATHER_STAGED=1" "ORBX_GATHER_STAGED=1 ORBX_NCCL_MAX_CTAS=16"; do
  i=$((i+1))
  echo "== $V"
  env $V timeout 120 $TR --master-port $((29530+i)) bench.py --gpus 8 --steps 20 --warmup 3 --value-only 2>gpurun_out/sw$i.err | tail -1
done
