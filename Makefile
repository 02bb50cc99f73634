# Convenience targets; everything is also reachable as plain python commands.
PY ?= python

.PHONY: build test test-gpu bench bench-ref sass clean

build:            ## nvcc -> fedrec-with-pytorchdistributed_b200/ops/libfedrec_b200.so (sm_100a, in-tree)
	$(PY) -c "import __graft_entry__ as g; g.build()"

test: build       ## CPU suite (gloo world_size 2-4 included)
	$(PY) -m pytest tests -q -m "not gpu"

test-gpu: build   ## on a B200
	$(PY) -m pytest tests -q -m gpu

bench: build      ## headline benchmark, one JSON line
	$(PY) bench.py --gpus 1 --steps 20 --warmup 5

bench-ref:        ## the unmodified upstream code on the same workload
	$(PY) bench.py --impl reference --gpus 1

sass: build       ## per-kernel SASS evidence into profiles/sass/
	$(PY) scripts/dump_sass.py

clean:
	rm -rf build fedrec-with-pytorchdistributed_b200/ops/libfedrec_b200.so
