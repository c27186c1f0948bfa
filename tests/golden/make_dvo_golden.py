"""Generates tests/golden/dvo_golden.npz by EXECUTING the reference's own Python for the pieces of
the DVO estimator that import here (numpy/scipy only):
    tadataka/vo/dvo/jacobian.py   (run by path: importing the package pulls in skimage)
    tadataka/robust/weights.py, tadataka/math.py, tadataka/se3.py
Run in the build container only (needs /root/reference):
    python tests/golden/make_dvo_golden.py
"""
import os
import runpy
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")
import tadataka.math as ref_math          # noqa: E402
import tadataka.robust.weights as ref_w   # noqa: E402
import tadataka.se3 as ref_se3            # noqa: E402

jac = runpy.run_path("/root/reference/tadataka/vo/dvo/jacobian.py")
rng = np.random.default_rng(20)
n = 200
P = np.column_stack((rng.normal(size=n) * 30, rng.normal(size=n) * 20, rng.uniform(50, 400, n)))
didx, didy = rng.normal(size=n) * 0.05, rng.normal(size=n) * 0.05
focal = np.array([517.3, 516.5])
J = jac["calc_jacobian"](focal, didx, didy, P)
image = rng.random((13, 17))
DX, DY = jac["calc_image_gradient"](image)
r = rng.normal(size=500) * 0.2
r[::25] *= 20          # outliers
w_huber, w_student, w_tukey = (ref_w.compute_weights_huber(r), ref_w.compute_weights_student_t(r),
                               ref_w.compute_weights_tukey(r))
A, b, w = rng.normal(size=(n, 6)), rng.normal(size=n), rng.uniform(0.1, 2.0, n)
x_plain, x_weighted = ref_math.solve_linear_equation(A, b), ref_math.solve_linear_equation(A, b, w)
xis = np.vstack([rng.normal(size=(6, 6)) * 0.1, np.array([[0.3, -0.2, 0.1, 0., 0., 0.]])])
ts = np.array([ref_se3.exp_se3_t_(xi) for xi in xis])
Gs = np.array([ref_se3.exp_se3(xi) for xi in xis])
np.savez_compressed(os.path.join(HERE, "dvo_golden.npz"), P=P, didx=didx, didy=didy, focal=focal, J=J,
                    image=image, DX=DX, DY=DY, r=r, w_huber=w_huber, w_student=w_student,
                    w_tukey=w_tukey, A=A, b=b, w=w, x_plain=x_plain, x_weighted=x_weighted, xis=xis,
                    ts=ts, Gs=Gs)
print("wrote dvo_golden.npz")
