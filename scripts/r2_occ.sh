cd $GRAFT_REPO_ROOT
python - <<'PY'
import kubectl_b200
from kubectl_b200 import _lib
L=_lib.lib()
import torch; torch.zeros(1).cuda()
for bn in (32,64,128):
    print(bn, [(s, L.kx_debug_max_active_clusters(bn, s)) for s in (1,2,4,8)])
PY
