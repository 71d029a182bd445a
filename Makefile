# Convenience targets for users coming from the reference (`make dllama dllama-api`): everything is built by build.py
# (nvcc for sm_100a + g++), in-tree, no install step.  `./dllama` and `./dllama-api` are the entry points.
PYTHON ?= python

.PHONY: all dllama dllama-api host cuda test test-gpu sanitize clean
all: dllama
dllama dllama-api:
	$(PYTHON) build.py
host:
	$(PYTHON) -c "import build; print(build.build_host())"
cuda:
	$(PYTHON) -c "import build; print(build.build_cuda())"
test: host
	$(PYTHON) -m pytest tests -q -m "not gpu"
test-gpu: dllama
	$(PYTHON) -m pytest tests -q -m gpu
sanitize:
	bash scripts/host_sanitizers.sh
clean:
	rm -rf build distributed-llama-multiusers_b200/lib/*.so
