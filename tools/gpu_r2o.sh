#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-r02o}
timeout 900 python -m pytest tests -m gpu -q -x -k "index_build" --durations=3 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $O/${TAG}_pytest.log
python - <<'PY'
import time, sys
sys.path.insert(0, ".")
import numpy as np
import vgaligner_b200 as vg
t0 = time.time(); ix = vg.Index.build_from_text(vg.synth_gfa(5000000, 100, 1), 15); t_host = time.time() - t0
t0 = time.time(); keys, koff, ts, te, ms = vg.index_build_device(ix, 15); t_dev = time.time() - t0
hs, he, ho = ix.table()
print("5 Mbp k=15: host build %.2f s (incl. GFA text + linearisation), device build call %.2f s, device time %.1f ms, keys %d, table %d, equal %s" %
      (t_host, t_dev, ms, len(keys), len(ts), np.array_equal(ts, hs) and np.array_equal(te, he) and np.array_equal(koff, ho) and np.array_equal(keys, np.ctypeslib.as_array(ix.view.key_kmer2bit, shape=(ix.view.n_keys,)))))
PY
