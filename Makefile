# Repo-level helper targets (the library itself is built by deepsdf_b200/csrc/Makefile or __graft_entry__.build()).
PY ?= python
.PHONY: build sass sanitize test-cpu

build:
	$(PY) -c "import __graft_entry__ as g; g.build()"

# per-kernel SASS listings of the hot kernels -> profiles/sass/ (no GPU needed)
sass: build
	$(PY) tools/dump_sass.py

# compute-sanitizer over the small fused paths (smoke step, many-tile forward, input-gradient and small-network consumers).
# Needs a B200; run through gpurun:  gpurun -- 'make sanitize'.  Logs -> gpurun_out/sanitize_*.log (copy into profiles/).
SAN_TESTS = tests/test_sanitize_targets.py
sanitize:
	mkdir -p gpurun_out
	for tool in memcheck racecheck synccheck; do \
	  timeout 900 compute-sanitizer --tool $$tool --print-limit 20 $(PY) -m pytest $(SAN_TESTS) -m gpu -x -q -s > gpurun_out/sanitize_$$tool.log 2>&1; \
	  echo "$$tool rc $$?" >> gpurun_out/sanitize_$$tool.log; tail -4 gpurun_out/sanitize_$$tool.log; \
	done

test-cpu:
	$(PY) -m pytest tests -x -q -m "not gpu"
