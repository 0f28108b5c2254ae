# Thin front end of the build (the real logic is parameter_server_b200/build.py: nvcc for sm_100a,
# g++ for the host runtime, everything in-tree).
PY ?= python

all:
	$(PY) -m parameter_server_b200.build

force:
	$(PY) -m parameter_server_b200.build --force

test: all
	$(PY) -m pytest tests -q -m "not gpu"

test-gpu: all
	$(PY) -m pytest tests -q -m gpu

bench: all
	$(PY) bench.py

sass: all
	$(PY) tools/sass_report.py

clean:
	rm -rf build parameter_server_b200/lib

.PHONY: all force test test-gpu bench sass clean tsan asan

# host runtime under sanitizers (build/selftest: native unit checks + in-process clusters + apps)
tsan:
	$(PY) tools/sanitize.py thread

asan:
	$(PY) tools/sanitize.py address
