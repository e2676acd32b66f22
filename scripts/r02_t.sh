#!/bin/bash
# run the given GPU test files, print the tail
timeout 900 python -m pytest "$@" -x -q 2>&1 | tail -30
