#!/bin/bash
timeout 300 python scripts/kernel_trace.py --workload cfg3 --steps 2 2>/dev/null | grep -E "^cfg3|k_gatv2_fwd" | cut -c1-150
