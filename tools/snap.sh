#!/bin/sh
# freeze a copy of the working tree under .snap/ (git- and history-ignored) so that a queued gpurun call keeps testing
# the state it was launched for while the main tree is edited and rebuilt
cd "$(dirname "$0")/.." && rm -rf .snap && mkdir .snap && \
tar --exclude=.git --exclude=gpurun_out --exclude=.snap --exclude=_obj --exclude=profiles --exclude=.scratch --exclude=__pycache__ --exclude=.pytest_cache -cf - . | tar -xf - -C .snap
