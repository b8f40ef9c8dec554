#!/bin/bash
for i in 1 2; do
  timeout 120 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print(d['value'], d['e2e']['value'], d['clocks'])"
done
timeout 120 python bench.py --steps 200 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print(d['value'], d['e2e']['value'], d['clocks'])"
