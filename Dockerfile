# gfedntm_b200 node image (server or client).  One container per federation node, like the reference's
# Dockerfile; the base image must provide CUDA 12.9 + PyTorch for sm_100a when the node owns a B200.
ARG BASE_IMAGE=nvcr.io/nvidia/pytorch:25.06-py3
FROM ${BASE_IMAGE}
ENV WORKDIR=/workspace PYTHONUNBUFFERED=1
WORKDIR /workspace
COPY . /workspace
RUN python -m pip install --no-cache-dir grpcio protobuf scikit-learn scipy pandas pyarrow && \
    python -m gfedntm_b200.ops.build
EXPOSE 50051-50060
ENTRYPOINT ["python", "main.py"]
