# Developer entry points (reference: top-level Makefile + tools/devcluster + .circleci).
PY ?= python

.PHONY: build test test-gpu bench bindings check tsan asan devcluster clean

build:            ## compile the native library for sm_100a (nvcc cross-compiles without a GPU)
	$(PY) -c "import __graft_entry__ as g; g.build()"

test:             ## CPU suite (multi-process paths run over gloo)
	$(PY) -m pytest tests/ -x -q -m "not gpu"

test-gpu:         ## needs a B200
	$(PY) -m pytest tests/ -x -q -m gpu

bench:            ## headline benchmark, one GPU
	$(PY) bench.py --gpus 1 --steps 100 --warmup 10

bindings:         ## regenerate common/api/bindings.py from the master's route table
	$(PY) -m determined_b200.common.api.generate_bindings

check: bindings   ## what CI runs before the tests
	git diff --exit-code determined_b200/common/api/bindings.py
	$(PY) scripts/cli_gaps.py --check
	$(PY) scripts/lint_names.py
	$(PY) scripts/gen_docs.py --check
	$(PY) -m compileall -q determined_b200 tests benchmarks examples bench.py __graft_entry__.py

tsan:             ## race check of the C++ host runtime (shared-memory collectives) under ThreadSanitizer
	g++ -std=c++17 -O1 -g -fsanitize=thread -pthread csrc/host/shm_collective.cpp csrc/host/shm_collective_test.cpp -lrt -o /tmp/detb200_shm_tsan
	TSAN_OPTIONS=exitcode=66 /tmp/detb200_shm_tsan 4 500

asan:             ## the C++ launcher under AddressSanitizer + UBSan (rank contract and kill-a-rank)
	$(PY) -m pytest tests/test_launch.py -q -k sanitizers

devcluster:       ## a master without own slots plus one agent, in the foreground of two background daemons
	$(PY) -m determined_b200.cli deploy local master-up --slots 0
	$(PY) -m determined_b200.cli deploy local agent-up
	@echo "dashboard: http://127.0.0.1:8080/   stop with: det deploy local agent-down && det deploy local master-down"

clean:
	rm -rf build determined_b200/_native/*.so determined_b200/_native/*.o .pytest_cache
