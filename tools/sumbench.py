import json,sys
for f in sys.argv[1:]:
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f,'ERR'); continue
    parts=f.split('/')[-1].replace('.json','').split('_')
    shape,fmt,M=parts[-3],parts[-2],int(parts[-1])
    E,k,H,I={'qwen3':(128,8,2048,768),'dsv3':(256,8,7168,2048),'gptoss':(128,4,2880,2880),'refbench':(256,8,7168,384)}[shape]
    fl=6*M*k*H*I/(d['ms_per_step']*1e-3)/1e12
    print(shape,fmt,M,'ms',round(d['ms_per_step'],4),'tok/s',int(d['value']),'TF',round(fl,1),'tc_frac',round(fl/1432.9,3),'hbm_frac',round(d['step_roofline'].get('hbm_gbs', d['step_roofline']['achieved'])/d['roofline'].get('peak',6450.0) if d['step_roofline']['bound']=='hbm' else d['step_roofline'].get('hbm_gbs',0)/6450.0,3),'bound',d['step_roofline']['bound'],d['config']['path'][:6],{k:round(v,1) for k,v in d['step_roofline']['stage_us'].items() if v>0},'e2e',int(d['e2e']['value']))
