# Convenience targets; the driver uses __graft_entry__.py, bench.py and pytest directly.
PY ?= python

build:            ## nvcc -> nanodiloco_b200/libnanodiloco_b200.so, gcc -> oracle/libnd_oracle.so
	$(PY) -c "import __graft_entry__ as g; g.build()"

test-cpu:         ## oracle vs golden vectors, host logic, ABI symbols, gloo world-size-2
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu:         ## parity through the C ABI (needs a B200)
	$(PY) -m pytest tests -x -q -m gpu

smoke:
	$(PY) -c "import __graft_entry__ as g; g.smoke()"

bench:
	$(PY) bench.py

golden:           ## regenerate tests/golden/*.npz from the reference (build container only)
	$(PY) tests/golden/make_golden.py

.PHONY: build test-cpu test-gpu smoke bench golden
