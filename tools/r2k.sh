#!/bin/bash
python tools/small_batch.py 8 2>&1 | tail -14
echo "--- dual halves"; OKRT_DUAL_MIN_FRAMES=2 python tools/small_batch.py 8 2>&1 | tail -7
echo "--- unsharded"; python tools/small_batch.py 1 2>&1 | tail -7
