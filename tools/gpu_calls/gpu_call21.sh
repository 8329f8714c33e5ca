#!/bin/bash
timeout 120 python tools/attn_prof.py 2>&1 | tail -2
