# Convenience targets; everything is plain Python underneath.
PY ?= python

build:            ## compile csrc/ for sm_100a into fed_tgan_b200/_native/libfedtgan_b200.so
	$(PY) __graft_entry__.py

test:             ## CPU test-suite (gloo / RPC included)
	$(PY) -m pytest tests -q -m "not gpu"

test-gpu:         ## GPU test-suite (needs a B200)
	$(PY) -m pytest tests -q -m gpu

bench:            ## headline metric on one GPU
	$(PY) bench.py --gpus 1 --steps 10 --warmup 3

bench-ref:        ## the unmodified reference on the same table
	$(PY) bench.py --impl reference --gpus 1 --steps 3 --warmup 1

.PHONY: build test test-gpu bench bench-ref
