"""bench.py against another build of the library: bench_with.py <libfrx.so> [bench.py flags]"""
import runpy, sys
from pathlib import Path
sys.path.insert(0, '/root/repo')
import flac_raster_b200._lib as L
L.LIB_PATH = Path(sys.argv[1]).resolve()
sys.argv = ['bench.py'] + sys.argv[2:]
runpy.run_path('/root/repo/bench.py', run_name='__main__')
