# Convenience wrapper; the build logic lives in build.py (nvcc, sm_100a only).
# Counterpart of the reference's cuda/Makefile and its README one-liners.
PY ?= python

all:
	$(PY) build.py

core:
	$(PY) build.py --no-torch

reference:
	$(PY) baseline/install_ref.py

test:
	$(PY) -m pytest tests -q -m "not gpu"

test-gpu:
	$(PY) -m pytest tests -q -m gpu

bench:
	$(PY) bench.py

clean:
	rm -rf build parallel-image-convolution-using-mpi-openmp-and-cuda_b200/lib

.PHONY: all core reference test test-gpu bench clean
