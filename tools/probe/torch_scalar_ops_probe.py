"""Which fp32 arithmetic does torch use on CUDA for tensor (op) python-scalar?  Prints which candidate matches."""
import torch
dev = torch.device("cuda")
torch.manual_seed(0)
x = (0.3 * torch.rand(1 << 20, device=dev) - 0.02)
xd = x.double()
b = 0.2508 - 0.02381
f32 = lambda v: torch.tensor(v, dtype=torch.float32, device=dev)
got = x / b
cands = {
    "true division by f32(b)": x / f32(b),
    "x * (1f / f32(b))": x * (f32(1.0) / f32(b)),
    "x * f32(1.0 / b)": x * f32(1.0 / b),
    "f32(double(x) / b)": (xd / b).float(),
}
for k, v in cands.items():
    print("div :", k, int((v != got).sum()))
got = x - 0.02381
print("sub : x - f32(c)", int(((x - f32(0.02381)) != got).sum()), " f32(double(x) - c)", int(((xd - 0.02381).float() != got).sum()))
got = 0.02381 + x
print("add : x + f32(c)", int(((x + f32(0.02381)) != got).sum()), " f32(double(x) + c)", int(((xd + 0.02381).float() != got).sum()))
got = b * x
print("mul : x * f32(b)", int(((x * f32(b)) != got).sum()), " f32(double(x) * b)", int(((xd * b).float() != got).sum()))
got = 0.02381 + b * x
print("expr: (x*f32(b)) + f32(c)", int((((x * f32(b)) + f32(0.02381)) != got).sum()))
