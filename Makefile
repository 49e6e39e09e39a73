# Convenience targets (everything is plain python underneath)
PY ?= python

.PHONY: build test test-gpu bench bench8 sass proto-check grpc-bench clean

build:            ## sm_100a native library, in-tree (no GPU needed)
	$(PY) -m dps_b200.build

test:             ## CPU tests: contract, codec, servicer numerics, gRPC e2e, failure handling, protocol model
	$(PY) -m pytest tests -q -m "not gpu"

test-gpu:         ## B200 tests: kernels vs fp32 PyTorch, engine, data pipeline, multi-GPU
	$(PY) -m pytest tests -q -m gpu

bench:            ## headline metric on one GPU (one JSON line)
	$(PY) bench.py --gpus 1 --steps 200 --warmup 10

bench8:
	$(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29500 \
		bench.py --gpus 8 --steps 200 --warmup 10

sass:             ## regenerate profiles/sass from the built library
	$(PY) bench_tools/dump_sass.py

proto-check:      ## ps.proto vs the programmatic descriptors
	$(PY) scripts/check_proto.py

grpc-bench:       ## BASELINE config 1 on CPU: this framework's gRPC path vs the reference arm
	$(PY) bench_tools/grpc_cpu_bench.py --workers 2 --steps 10 --with-reference

clean:
	rm -rf build dps_b200/_native/*.so out
