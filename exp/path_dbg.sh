#!/bin/sh
cd "$(dirname "$0")/.."
python exp/path_dbg.py 2>&1 | tail -22
RT_PATH_LEAF=1 RT_PATH_SERVE=1 python exp/path_dbg.py 2>&1 | tail -22
RT_PATH_LEAF=32 RT_PATH_SERVE=32 python exp/path_dbg.py 2>&1 | tail -22
