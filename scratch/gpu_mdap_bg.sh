#!/bin/bash
# background MDAP variants in the contract line's step: batch rows x grid cap
for rows in 0 1; do for cap in 8 4 2 1; do
JPLG_MDAP_ROWS8=$rows JPLG_MDAP_CTAS_PER_SM=$cap python bench.py --steps 300 --warmup 20 --no-cpu-baseline --no-e2e --no-capacity 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('rows8=$rows cap=$cap step', round(d['ms_per_step'],5), {k:round(x*1e3,1) for k,x in d['roofline']['kernel_ms_per_step'].items()}, d['bus_crc32'])"
done; done
