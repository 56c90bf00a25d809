import sys, os, math, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from music_style_transfer_ldm_b200 import ops
from gpu_util import nhwc
n, h, w, cin, cout, k = 16, 16, 128, 32, 128, 4
g = torch.Generator().manual_seed(1)
x = torch.randn((n, cin, h, w), generator=g).bfloat16()
wt = (torch.randn((cin, cout, k, k), generator=g) / math.sqrt(cin * k * k)).bfloat16().float()
xd = nhwc(x.float(), torch.bfloat16).cuda()
wp = ops.pack_weight(wt.cuda(), torch.bfloat16, transposed=True)
b = torch.randn(cout, generator=g).cuda()
st = torch.zeros(2 * cout, dtype=torch.float64, device="cuda")
y = ops.conv(xd, wp, b, ksize=4, stride=2, transposed=True, stats=st)
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()), float(st.sum()))
