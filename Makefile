.PHONY: build test test-gpu bench sass clean
build:
	python -m nanofed_b200.build
test:
	python -m pytest tests -q -m "not gpu"
test-gpu:
	python -m pytest tests -q -m gpu
bench:
	python bench.py
sass:
	python tools/dump_sass.py
clean:
	rm -rf build nanofed_b200/_native/*.so
