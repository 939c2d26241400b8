import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, math
from espnet_slurp_b200 import ops
torch.manual_seed(0)
for n_fft, hop in [(4, 2), (8, 4), (16, 8), (64, 32), (512, 128)]:
    N = max(4 * n_fft, 64)
    wav = torch.randn(1, N)
    nb = n_fft // 2 + 1
    eye = torch.eye(nb)
    f, fl = ops.logmel(wav.cuda(), torch.tensor([N], dtype=torch.int32).cuda(), eye.cuda(), n_fft=n_fft, hop=hop)
    x = torch.nn.functional.pad(wav[:, None], (n_fft // 2, n_fft // 2), mode="reflect")[:, 0]
    win = torch.hann_window(n_fft)
    T = N // hop + 1
    fr = torch.stack([x[0, t * hop: t * hop + n_fft] for t in range(T)])
    p = torch.fft.rfft(fr * win, dim=-1).abs() ** 2
    want = torch.clamp(p, min=1e-10).log()
    got = f[0].cpu()
    print(n_fft, "max err", (got - want).abs().max().item(), "frame1 got", got[1, :5].tolist(), "want", want[1, :5].tolist())
