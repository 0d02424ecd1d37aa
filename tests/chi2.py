"""TEST INFRASTRUCTURE: mi.chi2.ChiSquareTest (reference: src/python/python/chi2.py:5-323) restated on torch for
the emitter adapter (:530-567): histogram of sample_direction over a (phi, -cos theta) grid against pdf_direction
integrated per cell with the ires x ires trapezoid rule, low-expectation cells pooled as mi.math.chi2 does
(include/mitsuba/core/math.h:325-352), p-value from the chi^2 survival function."""
import numpy as np
import torch
from scipy.stats import chi2 as chi2_dist

SIN_OFFSET = 0.00775                                   # src/emitters/tests/test_sunsky.py:9


class SphericalDomain:
    """chi2.py:416-437: [phi, -cos(theta)] parameterisation of the unit sphere."""
    def bounds(self):
        return (-np.pi, -1.0), (np.pi, 1.0)

    def aspect(self):
        return 2

    def map_forward(self, px, py):
        cos_theta = -py
        sin_theta = torch.sqrt(torch.clamp(torch.addcmul(torch.ones_like(py), -cos_theta, cos_theta), min=0))
        return torch.stack([torch.cos(px) * sin_theta, torch.sin(px) * sin_theta, cos_theta])

    def map_backward(self, d):
        return torch.atan2(d[1], d[0]), -d[2]


class CroppedSphericalDomain(SphericalDomain):
    """test_sunsky.py:228-234: avoids the singularity at the north pole by SIN_OFFSET."""
    def bounds(self):
        cos_bound = float(np.sqrt(1 - SIN_OFFSET ** 2))
        return (-np.pi, -cos_bound), (np.pi, 1.0)


def pool_chi2(obs, exp, pool_threshold=5.0):
    """mi.math.chi2 on arrays already sorted by expected frequency (math.h:325-352)."""
    obs = np.asarray(obs, np.float64); exp = np.asarray(exp, np.float64)
    keep = ~((exp == 0) & (obs == 0))
    obs, exp = obs[keep], exp[keep]
    big = exp >= pool_threshold
    chsq = float((((obs[big] - exp[big]) ** 2) / exp[big]).sum())
    dof = int(big.sum())
    pooled_obs = pooled_exp = 0.0
    n_in = n_out = 0
    for o, e in zip(obs[~big], exp[~big]):              # sequential pooling, as the reference
        pooled_obs += o; pooled_exp += e; n_in += 1
        if pooled_exp > pool_threshold:
            chsq += (pooled_obs - pooled_exp) ** 2 / pooled_exp
            pooled_obs = pooled_exp = 0.0
            n_out += 1; dof += 1
    return chsq, dof - 1, n_in, n_out


class ChiSquareTest:
    def __init__(self, domain, sample_func, pdf_func, sample_count=1000000, res=101, ires=4, seed=0, batch=1 << 24,
                 device="cuda"):
        assert res > 0 and ires >= 2
        self.domain, self.sample_func, self.pdf_func = domain, sample_func, pdf_func
        self.sample_count, self.ires, self.seed, self.batch, self.device = sample_count, ires, seed, batch, device
        self.res = (max(int(res / domain.aspect()), 1), res)       # chi2.py:88-89: (x, y)
        self.bmin, self.bmax = domain.bounds()
        self.messages = []
        self.fail = False
        self.p_value = None

    def tabulate_histogram(self):
        rx, ry = self.res
        ext = (self.bmax[0] - self.bmin[0], self.bmax[1] - self.bmin[1])
        hist = torch.zeros(rx * ry, dtype=torch.float64, device=self.device)
        gen = torch.Generator(device=self.device); gen.manual_seed(self.seed)
        done = 0
        while done < self.sample_count:
            m = min(self.batch, self.sample_count - done)
            u = torch.rand((2, m), device=self.device, generator=gen)
            d = self.sample_func(u)
            x, y = self.domain.map_backward(d)
            eps = (ext[0] * 1e-4, ext[1] * 1e-4)
            ok = (x >= self.bmin[0] - eps[0]) & (x <= self.bmax[0] + eps[0]) & (y >= self.bmin[1] - eps[1]) & (y <= self.bmax[1] + eps[1])
            if not bool(ok.all()):
                self.messages.append("Encountered samples outside of the specified domain!")
                self.fail = True
            ix = ((x - self.bmin[0]) / ext[0] * rx).clamp(0, rx - 1).long()
            iy = ((y - self.bmin[1]) / ext[1] * ry).clamp(0, ry - 1).long()
            hist += torch.bincount(ix + iy * rx, minlength=rx * ry).double()
            done += m
        self.histogram = hist
        self.histogram_sum = float(hist.sum()) / self.sample_count

    def tabulate_pdf(self):
        rx, ry = self.res
        ires = self.ires
        ext = (self.bmax[0] - self.bmin[0], self.bmax[1] - self.bmin[1])
        cell = (ext[0] / rx, ext[1] / ry)
        spacing = (cell[0] / (ires - 1), cell[1] / (ires - 1))
        s = torch.arange(ires, device=self.device, dtype=torch.float32)
        off = (s + 1e-4) * (1 - 2e-4)                                # chi2.py:203
        w1 = torch.where((s == 0) | (s == ires - 1), 0.5, 1.0).double()
        pdf = torch.zeros(ry, rx, dtype=torch.float64, device=self.device)
        rows = 5
        for y0 in range(0, ry, rows):
            ys = torch.arange(y0, min(y0 + rows, ry), device=self.device, dtype=torch.float32)
            py = (self.bmin[1] + ys[:, None] * cell[1] + off[None, :] * spacing[1]).reshape(-1)      # (rows * ires)
            px = (self.bmin[0] + torch.arange(rx, device=self.device, dtype=torch.float32)[:, None] * cell[0]
                  + off[None, :] * spacing[0]).reshape(-1)                                          # (rx * ires)
            PY = py[:, None].expand(-1, px.numel()).reshape(-1)
            PX = px[None, :].expand(py.numel(), -1).reshape(-1)
            f = self.pdf_func(self.domain.map_forward(PX.float().contiguous(), PY.float().contiguous()).contiguous()).double()
            f = f.reshape(len(ys), ires, rx, ires) * w1[None, :, None, None] * w1[None, None, None, :]
            pdf[y0:y0 + len(ys)] = f.sum(dim=(1, 3)) * (spacing[0] * spacing[1] * self.sample_count)
        self.pdf = pdf.reshape(-1)
        self.pdf_sum = float(self.pdf.sum()) / self.sample_count
        if float(self.pdf.min()) < 0:
            self.messages.append("Failure: Encountered a cell with a negative PDF value"); self.fail = True
        if self.pdf_sum > 1.1:
            self.messages.append("Failure: PDF integrates to a value greater than 1.0"); self.fail = True

    def run(self, significance_level=0.01):
        self.tabulate_histogram()
        self.tabulate_pdf()
        pdf = self.pdf.cpu().numpy(); hist = self.histogram.cpu().numpy()
        order = np.argsort(pdf, kind="stable")
        chi2val, dof, pooled_in, pooled_out = pool_chi2(hist[order], pdf[order], 5)
        if dof < 1:
            self.messages.append("Failure: The number of degrees of freedom is too low!"); self.fail = True
        if np.any((pdf == 0) & (hist != 0)):
            self.messages.append("Failure: Found samples in a cell with expected frequency 0."); self.fail = True
        self.chi2val, self.dof = chi2val, dof
        self.p_value = float(chi2_dist.sf(chi2val, dof))
        self.messages.append(f"Histogram sum = {self.histogram_sum:f}, PDF sum = {self.pdf_sum:f}; "
                             f"Chi^2 statistic = {chi2val:f} (d.o.f = {dof}); pooled {pooled_in} -> {pooled_out}; p = {self.p_value:f}")
        return (not self.fail) and np.isfinite(self.p_value) and self.p_value >= significance_level
