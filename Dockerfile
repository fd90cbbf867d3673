# Blackwell needs CUDA >= 12.8; the devel image carries nvcc for csrc/ (sm_100a).
FROM pytorch/pytorch:2.7.0-cuda12.8-cudnn9-devel

RUN apt-get update && apt-get install -y --no-install-recommends git openssh-client tmux \
    && rm -rf /var/lib/apt/lists/*

WORKDIR /workspace
COPY requirements.txt .
RUN pip install --no-cache-dir -r requirements.txt
COPY . .
# nvcc cross-compiles for sm_100a; no GPU is needed at image build time
RUN python -m b200train.build
