python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python tools/tile_check.py
MPCB_DEBUG_TIMERS=1 python tools/prof_solve.py 2>&1 | tail -3
python tools/prof_solve.py 2>&1 | tail -2
PROF_B=512 python tools/prof_solve.py 2>&1 | tail -1
PROF_B=8192 python tools/prof_solve.py 2>&1 | tail -1
PROF_B=16384 python tools/prof_solve.py 2>&1 | tail -1
