# Developer shortcuts (the reference's Makefile is empty although its CONTRIBUTING mentions `make lint`).
PY ?= python

.PHONY: build test test-gpu bench smoke sass clean

build:
	$(PY) -m ddlb_b200._native.build --force

test:
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu:
	$(PY) -m pytest tests -x -q -m gpu

bench:
	$(PY) bench.py

smoke:
	$(PY) -c "import __graft_entry__ as g; g.build(); g.smoke()"

sass: build
	cuobjdump -sass ddlb_b200/_native/libddlb_b200.so | grep -oE "UTC[A-Z]*MMA[A-Z0-9.]*|UTMA(LDG|STG)[A-Z0-9.]*|LDTM[A-Z0-9.]*|UBLKCP[A-Z0-9.]*" | sort | uniq -c

clean:
	rm -f ddlb_b200/_native/libddlb_b200.so
	rm -rf results gpurun_out .pytest_cache
