import sys, json
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from applied_image_processing_b200 import _cabi, localized_style_transfer as lst
from tests import fullsize
out = {}
for name in ("fs_4k_U", "fs_4k_N", "fs_1080p_U"):
    spec = fullsize.CASES[name]
    c, s, m = fullsize.inputs(spec)
    plan = lst.Plan(torch.device("cuda", 0), spec["H"], spec["W"], spec["sh"], spec["sw"], slots=1)
    dc, ds, dm = (torch.from_numpy(x).cuda() for x in (c, s, m))
    o = torch.empty_like(dc)
    _cabi.check(_cabi.lib().rct_transfer_async(plan.handle, dc.data_ptr(), c.size, ds.data_ptr(), s.size, dm.data_ptr(), m.size, o.data_ptr(), c.size, 1, 0, 0, torch.cuda.current_stream().cuda_stream))
    out[name] = plan.debug_keytab()
print(json.dumps(out))
