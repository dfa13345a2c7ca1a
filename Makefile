# Build / test / bench shortcuts (the reference builds a fat jar with `gradle build`, build.gradle:19-30).
.PHONY: build test test-gpu bench clean mock-data
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
