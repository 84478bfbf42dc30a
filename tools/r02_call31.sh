#!/bin/bash
timeout 120 python tools/attn_trace_tmp.py 2>&1 | tail -8
