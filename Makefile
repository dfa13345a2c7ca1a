# Build / test / bench shortcuts (the reference builds a fat jar with `gradle build`, build.gradle:19-30).
.PHONY: build test test-gpu bench clean mock-data tsan asan sass
build:
	./build.sh
test: build
	python -m pytest tests -x -q -m "not gpu"
test-gpu: build
	python -m pytest tests -x -q -m gpu
bench: build
	python bench.py
mock-data:
	python tools/make_mock_data.py
clean:
	rm -rf build parameter-server-architecture-on-apache-kafka_b200/ops/libpsb200.so
# ThreadSanitizer run of the C++ host runtime (producer thread vs consumer / slot-assignment threads); host only
tsan:
	mkdir -p build
	g++ -std=c++17 -O1 -g -fsanitize=thread -fPIE -pie -x c++ -I/usr/local/cuda/include csrc/host_runtime.cpp csrc/host_producer.cpp csrc/host_pack.cpp \
	    tools/tsan_host_runtime.cpp tools/tsan_stubs.cpp -o build/tsan_host_runtime -lpthread
	TSAN_OPTIONS="halt_on_error=1 second_deadlock_stack=1" ./build/tsan_host_runtime
# AddressSanitizer + UBSan run of the host packers and the CSV tokenizer on buffers of exactly the documented size,
# once with the SSE2 paths and once with the portable ones; host only
asan:
	mkdir -p build
	for defs in "" "-DPSB_NO_SIMD"; do \
	  g++ -std=c++17 -O1 -g $$defs -fsanitize=address,undefined,float-cast-overflow,float-divide-by-zero -fno-sanitize-recover=all -x c++ \
	      -I/usr/local/cuda/include csrc/host_runtime.cpp csrc/host_producer.cpp csrc/host_pack.cpp tools/asan_packers.cpp tools/tsan_stubs.cpp \
	      -o build/asan_packers -lpthread && ./build/asan_packers || exit 1; \
	done
# full SASS listing of every kernel + mnemonic summary under profiles/sass (no GPU needed)
sass: build
	python tools/dump_sass.py
