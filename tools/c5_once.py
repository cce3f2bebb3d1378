import json, os, sys
sys.path.insert(0, '/root/repo')
import bench
from atooms_b200 import capi
lib = capi.load()
r = bench.c5_strong(lib, 0, 1, 0)
print(json.dumps({'env': os.environ.get('B2_LIST_WINDOWED'), 'value': r['value'], 'ms': r['ms_per_md_step'], 'refreshes': r['refreshes']}))
