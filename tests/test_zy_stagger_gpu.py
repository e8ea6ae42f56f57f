"""The Yee staggering of the gather on the CUDA path (tiled kernels, TMA-staged tiles in 2D/3D): linear fields sampled at
each component's own Yee locations must be gathered exactly -- the same check as
tests/test_oracle.py::test_gather_staggering_linear_fields, through the product library (tolerance 1e-11 of the field
value; a wrong stagger costs 1e-2)."""
import pytest

from helpers import SHAPE_NAMES
from test_oracle import check_linear_gather, check_particle_bcs


@pytest.mark.gpu
@pytest.mark.parametrize("ndims", [1, 2, 3])
@pytest.mark.parametrize("shape", SHAPE_NAMES)
def test_gather_staggering_linear_fields_gpu(product_abi, ndims, shape):
    check_linear_gather(product_abi, ndims, shape, tol=1e-11)


@pytest.mark.gpu
@pytest.mark.parametrize("ndims", [1, 2, 3])
def test_particle_bcs_known_answers_gpu(product_abi, ndims):
    """placed particles through eb200_particle_bcs: wrap, mirror, removal beyond the outer edge, < on the min side and
    >= on the max side -- bit for bit the known answers of tests/test_oracle.py::test_particle_bcs_known_answers"""
    check_particle_bcs(product_abi, ndims)
