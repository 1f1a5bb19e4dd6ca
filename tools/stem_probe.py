"""Probe: which cuDNN route is fastest for the DLA stem (7x7, 3->16 at 1024x2048) and level0/level1 convs in bf16."""
import torch, torch.nn.functional as F
torch.backends.cudnn.benchmark = True
dev = "cuda"
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1000
H, W = 1024, 2048
for cin in (3, 4, 8, 16):
    for dt in (torch.bfloat16, torch.float16):
        x = torch.randn(1, cin, H, W, device=dev, dtype=dt).contiguous(memory_format=torch.channels_last)
        w = torch.randn(16, cin, 7, 7, device=dev, dtype=dt).contiguous(memory_format=torch.channels_last)
        b = torch.zeros(16, device=dev, dtype=dt)
        try:
            print(f"stem cin={cin} {dt} plain conv      : {t(lambda: F.conv2d(x, w, None, 1, 3)):8.1f} us")
            print(f"stem cin={cin} {dt} conv+bias       : {t(lambda: F.conv2d(x, w, b, 1, 3)):8.1f} us")
            print(f"stem cin={cin} {dt} cudnn conv_relu : {t(lambda: torch.cudnn_convolution_relu(x, w, b, (1,1), (3,3), (1,1), 1)):8.1f} us")
        except Exception as e:
            print("err", cin, dt, str(e)[:80])
x = torch.randn(1, 16, H, W, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
w = torch.randn(16, 16, 3, 3, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
b = torch.zeros(16, device=dev, dtype=torch.bfloat16)
print(f"level0 16->16 3x3 plain      : {t(lambda: F.conv2d(x, w, None, 1, 1)):8.1f} us")
print(f"level0 16->16 3x3 conv_relu  : {t(lambda: torch.cudnn_convolution_relu(x, w, b, (1,1), (1,1), (1,1), 1)):8.1f} us")
w2 = torch.randn(32, 16, 3, 3, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
b2 = torch.zeros(32, device=dev, dtype=torch.bfloat16)
print(f"level1 16->32 3x3 s2 plain   : {t(lambda: F.conv2d(x, w2, None, 2, 1)):8.1f} us")
print(f"level1 16->32 3x3 s2 conv_relu: {t(lambda: torch.cudnn_convolution_relu(x, w2, b2, (2,2), (1,1), (1,1), 1)):8.1f} us")
xf = torch.randn(1, 3, H, W, device=dev).contiguous(memory_format=torch.channels_last)
wf = torch.randn(16, 3, 7, 7, device=dev).contiguous(memory_format=torch.channels_last)
print(f"stem fp32 cin=3 plain        : {t(lambda: F.conv2d(xf, wf, None, 1, 3)):8.1f} us")
# max pool
for c, h in ((32, 512), (64, 256), (128, 128)):
    xm = torch.randn(1, c, h, 2 * h, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    print(f"maxpool2 c={c} h={h}: {t(lambda: F.max_pool2d(xm, 2, 2)):8.1f} us")
