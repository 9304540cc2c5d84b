"""One-line digest of a bench.py JSON line (development aid).  Usage: python scripts/show_bench.py FILE"""
import json,sys
j=json.loads([l for l in open(sys.argv[1]).read().strip().splitlines() if l.startswith('{')][-1])
print(sys.argv[1], 'value',round(j["value"],1),'ms',round(j["ms_per_step"],3),'dev',round(j["device_ms_per_step"],3),'prove_ms',round(j["prove_ms"],3),'e2e',round(j["e2e"]["value"],1), 's22', j.get('s22') and (round(j['s22']['prove_ms_device'],2), round(j['s22']['prove_ms_wall'],2)), 'wm', j.get('witness_map') and [round(v['ms'],3) for k,v in j['witness_map'].items() if k!='note'])
