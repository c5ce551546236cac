# rabit-b200: thin make front end over build.py (in-tree build of lib/*.so and build/bin/*)
.PHONY: all host test gputest clean lint sanitize
all:
	python build.py
host:
	python build.py --no-cuda
test: all
	python -m pytest tests -x -q -m "not gpu"
gputest: all
	python -m pytest tests -x -q -m gpu
clean:
	python build.py --clean
lint:
	python scripts/lint.py
sanitize:
	python scripts/sanitize.py address && python scripts/sanitize.py thread
