# Build / test / package targets (reference: make-dist.sh + the Maven modules' Makefile.common).
PY ?= python

.PHONY: build test test-gpu bench dist docs sanitize clean

build:            ## nvcc (sm_100a) + g++ -> bigdl_b200/lib/*.so
	$(PY) -m bigdl_b200.build

test: build       ## CPU suite (multi-process paths on gloo)
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu: build   ## kernels vs fp32 PyTorch references, needs a B200
	$(PY) -m pytest tests -x -q -m gpu

bench: build
	$(PY) bench.py --gpus 1

dist: build       ## wheel with the native libraries + a dist/ tree like the reference's (lib, conf, scripts)
	$(PY) setup.py -q bdist_wheel
	mkdir -p dist/lib dist/conf dist/bin
	cp bigdl_b200/lib/*.so dist/lib/
	cp conf/*.conf dist/conf/
	cp scripts/*.sh dist/bin/

docs:
	$(PY) docs/gen_api_lists.py

sanitize: build   ## compute-sanitizer over the kernel tests (memcheck; racecheck/synccheck: TOOL=racecheck)
	compute-sanitizer --tool $${TOOL:-memcheck} $(PY) -m pytest tests/test_kernels_gpu.py -x -q -m gpu

clean:
	rm -rf build dist *.egg-info bigdl_b200/lib/*.so bigdl_b200/lib/.*.sha256
