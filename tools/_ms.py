import json,sys
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(sys.argv[1], 'ms/step', round(d['ms_per_step'],3), 'img/s', round(d['value']), 'launches', d.get('gpu_launches'))
