"""Distance -> similarity kernels (reference src/similarity.py:20-76)."""
import numpy as np


class SimilarityKernel(object):
    def name(self):
        return ''

    def shortname(self):
        return ''

    def name_suffix(self):
        return ''

    def dist_to_sim_np(self, dist, *unused):
        raise NotImplementedError()

    # the reference's second twin builds a TF op; on this path the device epilogue
    # (csrc/sgnn_common.cuh: SGNN_ACT_SIM_KERNEL) plays that role, so both names map to numpy
    def dist_to_sim_tf(self, dist, *unused):
        return self.dist_to_sim_np(dist)


class IdentityKernel(SimilarityKernel):
    def dist_to_sim_np(self, dist, *unused):
        return dist


class GaussianKernel(SimilarityKernel):
    def __init__(self, yeta):
        self.yeta = yeta

    def name(self):
        return 'Gaussian_yeta={:g}'.format(self.yeta)

    def shortname(self):
        return 'g_{:.2e}'.format(self.yeta)

    def dist_to_sim_np(self, dist, *unused):
        return np.exp(-self.yeta * np.square(dist))


def create_sim_kernel(kernel_name, yeta=None):
    if kernel_name == 'identity':
        return IdentityKernel()
    if kernel_name == 'gaussian':
        return GaussianKernel(yeta)
    raise RuntimeError('Unknown sim kernel {}'.format(kernel_name))
