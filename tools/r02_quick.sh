#!/bin/bash
# quick GPU check: full parity suite, smoke, one bench line
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) 2>&1 | tail -6
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" ) 2>&1 | tail -3
