#!/usr/bin/env python
"""print the per-kernel table of bench.py JSON lines:  python tools/bench_show.py file.json [...]"""
import json
import sys

for path in sys.argv[1:]:
    for l in open(path):
        l = l.strip()
        if not l.startswith('{'):
            continue
        d = json.loads(l)
        print('==', path, 'ms_per_step %.2f' % d['ms_per_step'], {k: round(v, 2) for k, v in d.get('stages_ms', {}).items()},
              'e2e', d.get('e2e', {}).get('ms_per_step'))
        print('   rounds', d.get('lms_sort_rounds'), 'tied', d.get('lms_sort_tied'), 'launches', d.get('gpu_launches'))
        for k, v in d.get('kernels', {}).items():
            print('  %-28s %8.3f ms  %5.1f launches  frac %s' % (k, v['ms_per_step'], v['launches_per_step'],
                                                                 ('%.3f' % v['frac']) if 'frac' in v else '-'))
            for ck, cv in v.get('classes', {}).items():
                print('      %-28s %8.3f ms %5.1f' % (ck, cv['ms_per_step'], cv['launches_per_step']))
