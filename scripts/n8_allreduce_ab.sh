run() { tag=$1; shift; env "$@" timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus 8 --steps 40 --warmup 5 --skip-train --skip-e2e --skip-cpu-baseline --no-clock-probe > gpurun_out/n8_$tag.json 2> gpurun_out/n8_$tag.err; echo $tag rc=$?; }
run nccl X=1
run mc PSGD_PEER_ALLREDUCE=1
run pull PSGD_PEER_ALLREDUCE=1 PSGD_PEER_MULTICAST=0
run nvls NCCL_ALGO=NVLS
run ll NCCL_PROTO=LL
run tree NCCL_ALGO=Tree
