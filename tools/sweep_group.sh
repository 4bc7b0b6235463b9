#!/bin/bash
# device-resident cfg 2 throughput against the number of frames per kernel launch group
for g in 1 2 3 4 6 8 16 32; do
  python bench.py --no-cpu --no-e2e --group $g 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('group', d['config'].get('group_frames', '?'), round(d['value']), round(d['roofline']['pipeline']['frac'], 4))"
done
