# One image for every role, like the reference's Dockerfile (python:3.10-slim + uvicorn server:app), but for B200 nodes:
# CUDA runtime base, the library built for sm_100a at image build time (nvcc cross-compiles without a GPU).
FROM nvcr.io/nvidia/pytorch:25.03-py3
WORKDIR /app
COPY include ./include
COPY llm_sharding_demo_b200 ./llm_sharding_demo_b200
RUN pip install --no-cache-dir fastapi "uvicorn[standard]" requests \
 && python -m llm_sharding_demo_b200.build --force
ENV MODEL_ID=gpt2 SHARD_ROLE=coordinator SPLIT_AT=6 SHARD_PORT=5000
EXPOSE 5000
# single-GPU roles (a | b | coordinator): the reference's own entry point shape (Dockerfile:19 `uvicorn server:app`)
#   docker run --gpus 1 -e SHARD_ROLE=a ... IMAGE
# N-GPU pipeline pod (k8s/pipeline-pod.yaml):
#   docker run --gpus all IMAGE torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 -m llm_sharding_demo_b200.pod
CMD ["uvicorn", "llm_sharding_demo_b200.server:app", "--host", "0.0.0.0", "--port", "5000"]
