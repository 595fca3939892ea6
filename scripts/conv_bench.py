"""conv_igemm_tcgen05 forward vs cuDNN (F.conv2d, channels_last bf16) on ResNet-50 / ResNet-18 layer shapes."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.nn.functional as F

from hyperion_b200.ops import conv as convops
from gemm_layer_bench import timeit

torch.backends.cudnn.benchmark = True
dt = torch.bfloat16
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
for (C, H, K, R, stride, pad) in [(64, 56, 64, 3, 1, 1), (128, 28, 128, 3, 1, 1), (256, 14, 256, 3, 1, 1), (512, 7, 512, 3, 1, 1),
                                  (128, 56, 128, 3, 2, 1), (256, 28, 256, 3, 2, 1), (256, 56, 512, 1, 2, 0), (64, 56, 256, 1, 1, 0)]:
    x = torch.randn(B, C, H, H, device="cuda", dtype=dt).contiguous(memory_format=torch.channels_last)
    w = torch.randn(K, C, R, R, device="cuda", dtype=dt).contiguous(memory_format=torch.channels_last)
    Ho = (H + 2 * pad - R) // stride + 1
    fl = 2.0 * B * Ho * Ho * K * C * R * R / 1e9
    t_h = timeit(lambda: convops.conv2d_igemm(x, w, stride, pad))
    t_c = timeit(lambda: F.conv2d(x, w, None, stride, pad))
    dy = torch.randn(B, K, Ho, Ho, device="cuda", dtype=dt).contiguous(memory_format=torch.channels_last)
    t_wh = timeit(lambda: convops.conv2d_wgrad_igemm(x, dy, w.shape, stride, pad))
    t_wc = timeit(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, [stride] * 2, [pad] * 2, [1, 1], False, [0, 0], 1, [False, True, False]))
    t_dc = timeit(lambda: torch.ops.aten.convolution_backward(dy, x, w, None, [stride] * 2, [pad] * 2, [1, 1], False, [0, 0], 1, [True, False, False]))
    print(f"B{B} C{C} {H}x{H} -> K{K} {R}x{R}/s{stride}: fwd igemm {t_h * 1e3:6.1f} us cudnn {t_c * 1e3:6.1f} us | "
          f"wgrad igemm(+zero,+round) {t_wh * 1e3:6.1f} us cudnn {t_wc * 1e3:6.1f} us | dgrad cudnn {t_dc * 1e3:6.1f} us", flush=True)
