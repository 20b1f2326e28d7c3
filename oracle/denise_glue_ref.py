"""CPU restatement of what the reference does around DENISE's gradient.  TEST INFRASTRUCTURE ONLY.

* `gradient_post`: models/networks.py:9194-9240 line by line in numpy (this part of the path IS in the reference, so this
  oracle is pinned by the lines it restates); the flipud pairs (:9194-9196 undo :8964-8978) cancel for row-0-at-the-surface
  arrays and are left out, as in the product.
* `lowpass`: the frequency stage of `d.add_fwi_stage(fc_low=0.0, fc_high=10.0)` (:9147).  DENISE is external: the filter is
  restated as scipy's Butterworth second-order sections applied forward and then backward with zero initial state
  (PARITY UNPINNED for DENISE's exact filter order / phase).
"""
import numpy as np


def gradient_post(vp_grad, vs_grad, rho_grad, vpst, vsst, rhost, water_rows=24):
    vp_grad, vs_grad, rho_grad = (np.array(g, dtype=np.float32, copy=True) for g in (vp_grad, vs_grad, rho_grad))
    g1 = np.arange(np.shape(rho_grad)[0])                       # :9198
    g1 = g1 ** 1.5                                              # :9199
    ss = rho_grad * 0                                           # :9200
    for i in range(np.shape(rho_grad)[1]):                      # :9201-9202
        ss[:, i] = g1
    rho_grad = rho_grad * ss                                    # :9203
    vp_grad[0:water_rows, :] = 0.0                              # :9205
    vs_grad[0:water_rows, :] = 0.0                              # :9206
    rho_grad[0:water_rows, :] = 0.0                             # :9207
    r1 = np.max(np.abs(vpst)) / np.max(np.abs(vp_grad))         # :9220
    vp_grad = 1.0 * vp_grad.astype(np.float32) * r1             # :9221-9223
    r2 = np.max(np.abs(vsst)) / np.max(np.abs(vs_grad))         # :9226
    vs_grad = 1.0 * vs_grad.astype(np.float32) * r2             # :9227-9229
    r3 = np.max(np.abs(rhost)) / np.max(np.abs(rho_grad))       # :9231
    rho_grad = 0.5 * rho_grad.astype(np.float32) * r3           # :9232-9234
    rho_grad = rho_grad + 1e-10                                 # :9235
    return vp_grad, vs_grad, rho_grad


def lowpass(x, dt, fc_high, order=6, zero_phase=True):
    """x [nt, ...]: Butterworth low-pass along axis 0, zero initial state."""
    from scipy.signal import butter, sosfilt
    sos = butter(order, fc_high, btype="low", fs=1.0 / dt, output="sos")
    y = sosfilt(sos, np.asarray(x, dtype=np.float64), axis=0)
    if zero_phase:
        y = sosfilt(sos, y[::-1], axis=0)[::-1]
    return y


def lowpass_matrix(nt, dt, fc_high, order=6, zero_phase=True):
    """The filter as a dense [nt, nt] matrix (columns = responses to unit impulses) for small autograd checks."""
    return lowpass(np.eye(nt), dt, fc_high, order, zero_phase)
