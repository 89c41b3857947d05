#!/bin/bash
# dev helper: rebuild the library (abort on failure), then run a command on the GPU box
set -e
cd /root/repo
python -c "from network_distributed_pytorch_b200.build import build; print(build())"
exec /usr/local/graft/bin/gpurun --timeout "${GPU_TIMEOUT:-1500}" -- "$@"
