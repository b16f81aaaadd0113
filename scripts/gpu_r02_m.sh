#!/bin/bash
# allocator effect on the fixed list: expandable segments on/off, 1 and 2 streams
set -x
mkdir -p gpurun_out
for conf in default expandable; do for st in 1 2; do
  if [ $conf = expandable ]; then export PYTORCH_CUDA_ALLOC_CONF=expandable_segments:True; else unset PYTORCH_CUDA_ALLOC_CONF; fi
  timeout 600 python bench.py --scenes 312 --pool 8 --streams $st --no-cpu-baseline > gpurun_out/m_list_${conf}_s$st.json 2> gpurun_out/m_list_${conf}_s$st.err
  python -c "
import json; l=json.load(open('gpurun_out/m_list_${conf}_s$st.json')); print('$conf', $st, l['value'], l['e2e']['value'], l['seconds'], l['clocks']['sm_mhz'])"
done; done
