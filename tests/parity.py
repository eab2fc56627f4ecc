"""Shared tolerance protocol: per-tensor rel-L2 against the reference's fp64
result; fp32 path <= max(1e-5, 4 x the reference's own fp32-vs-fp64 error of that tensor)."""
from conftest import rel_l2

FP32_TOL = 1e-5
TC_TOL = 2e-3


def check_case(G, out, grads, gin, floor=FP32_TOL, noise_mult=4.0):
    fails, worst = [], 0.0

    def one(label, got, want, noise):
        nonlocal worst
        if got is None:
            fails.append(f"{label}: missing")
            return
        e = rel_l2(got, want)
        worst = max(worst, e)
        tol = max(floor, noise_mult * noise)
        if not e < tol:
            fails.append(f"{label}: rel-L2 {e:.3e} > tol {tol:.3e} (reference fp32 noise {noise:.2e})")

    one("out", out, G.out, G.noise("out"))
    for name, g in G.grads.items():
        one("grad " + name, grads.get(name), g, G.noise("g", name))
    for name, g in G.gin.items():
        if name == "adjacency":
            continue
        one("dinput " + name, gin.get(name), g, G.noise("gin", name))
    return fails, worst
