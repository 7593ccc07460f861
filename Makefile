# Convenience targets; the build itself is deepfm_b200/build.py (nvcc -gencode arch=compute_100a,code=sm_100a + g++, in-tree .so files).
PY ?= python

.PHONY: build test test-gpu bench sass clean

build:
	$(PY) -m deepfm_b200.build

test: build
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu: build
	$(PY) -m pytest tests -x -q -m gpu

bench: build
	$(PY) bench.py --gpus 1

sass: build
	$(PY) scripts/sass_evidence.py > profiles/sass_evidence.md

clean:
	rm -rf build deepfm-tensorflow-distributed-training-on-sagemaker_b200/ops/_lib deepfm-tensorflow-distributed-training-on-sagemaker_b200/data/_lib
