import json,sys
l=[x for x in open(sys.argv[1]) if x.startswith('{')][-1]
d=json.loads(l)
print('value',round(d['value']), 'ms/step',round(d['ms_per_step'],3), 'e2e',round(d['e2e']['value']), 'clk', d['clocks'])
for k,v in d['roofline']['per_kernel'].items(): print('  ',k, round(v['ms'],3), v['gbs'] and round(v['gbs']))
