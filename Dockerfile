# Build + test image.  The CUDA 12.9 toolkit is needed for sm_100a (B200); the library itself needs no GPU to BUILD.
FROM nvidia/cuda:12.9.0-devel-ubuntu24.04
RUN apt-get update && apt-get install -y --no-install-recommends python3 python3-pip python3-venv git && rm -rf /var/lib/apt/lists/*
WORKDIR /opt/myfyp_b200
COPY . .
RUN python3 -m venv /opt/venv && /opt/venv/bin/pip install --upgrade pip && \
    /opt/venv/bin/pip install torch numpy grpcio pytest pytest-timeout psutil nvidia-ml-py pandas pyarrow && \
    /opt/venv/bin/pip install -e . && \
    /opt/venv/bin/python -m myfyp_b200.build
ENV PATH=/opt/venv/bin:$PATH
# CPU-side suite by default; on a B200 host: docker run --gpus all ... python -m pytest tests -q -m gpu
CMD ["python", "-m", "pytest", "tests", "-q", "-m", "not gpu"]
