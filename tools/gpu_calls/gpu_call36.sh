#!/bin/bash
export SXB_ATTN_FWD128=2
for m in 0 4 5 6 7 8; do SXB_ATTN_DEBUG=$m timeout 100 python tools/attn_debug.py 2>&1 | tail -1 | cut -c1-120; done
