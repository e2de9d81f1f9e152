# B200 (sm_100a) image: CUDA 12.9 toolchain + PyTorch cu128 wheels + the in-tree kernel library.
FROM nvidia/cuda:12.9.0-devel-ubuntu24.04
ENV DEBIAN_FRONTEND=noninteractive PIP_BREAK_SYSTEM_PACKAGES=1
RUN apt-get update && apt-get install -y --no-install-recommends python3 python3-pip python3-dev git g++ ninja-build \
    && rm -rf /var/lib/apt/lists/* && ln -sf /usr/bin/python3 /usr/bin/python
RUN python -m pip install --index-url https://download.pytorch.org/whl/cu128 "torch>=2.7"
WORKDIR /workspace/llm-training
COPY . .
RUN python build_kernels.py && python -m pip install --no-build-isolation -e .
ENTRYPOINT ["llm-training"]
