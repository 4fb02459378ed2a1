# Convenience targets; everything is plain python underneath.
PY ?= python
.PHONY: build test test-gpu bench bench-ref sass-report clean
build:
	$(PY) -m distributed_llm_code_samples_b200.build_native
test: build
	$(PY) -m pytest tests -x -q -m "not gpu"
test-gpu: build
	$(PY) -m pytest tests -x -q -m gpu
sass-report: build   # regenerate profiles/sass and profiles/ptxas (no GPU needed)
	$(PY) tools/sass_report.py
bench: build
	$(PY) bench.py --gpus 1
bench-ref:
	$(PY) bench.py --impl reference --gpus 1
clean:
	rm -rf distributed_llm_code_samples_b200/_native
