.PHONY: build test test-gpu bench sass clean docs release lint
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
docs:
	python -m sphinx -b html docs/source docs/_build/html
lint:
	python -m compileall -q nanofed_b200 tools tests bench.py
release:
	@test -n "$(VERSION)" || (echo "usage: make release VERSION=X.Y.Z"; exit 1)
	scripts/release.sh $(VERSION)
