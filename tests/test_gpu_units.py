"""GPU: every device function of the hot path, run in a one-thread kernel through the C-ABI
(orb_unit_*), against the golden vectors generated from the reference source."""
import numpy as np
import pytest

import _parity as P
from optix_renderer_b200 import scenes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def r(renderer_factory):
    return renderer_factory(scenes.scene_floor_blocker(8, 8))


@pytest.fixture(scope="module")
def ora(restatement):
    restatement.load_scene(scenes.scene_floor_blocker(8, 8))
    return restatement


def test_rng_bit_exact(r):
    P.check_rng(r, P.units())


def test_clip_and_edge_sum(r, ora):
    P.check_clip(r, ora, P.units())


def test_scalar_units(r):
    P.check_scalar_units(r, P.units())


def test_bsdf(r):
    P.check_bsdf(r, P.units())


def test_ltc_pixel_survey_kat(renderer_factory):
    """ltcDirectLighingBaseline for one shading point under a unit quad."""
    a, b, c, d = (-.5, -.5, 1), (.5, -.5, 1), (.5, .5, 1), (-.5, .5, 1)
    light = scenes.quad_mesh(a, b, c, d, (0, 0, -1), diffuse=(0, 0, 0), alpha=1.0, emit=(1, 1, 1), is_light=True,
                             material_id=0)
    cam = scenes.camera_from_lookat((1, 0.5, 1.5), (0, 0, 0), (0, 0, 1), 0.66, 8, 8)
    sc = scenes.finish([light], cam, 8, 8)
    r = renderer_factory(sc)
    want = {0.01: (1.18095469, 1.13571608, 1.09047723), 0.25: (0.669329166, 0.624090433, 0.5788517),
            0.50: (0.379051447, 0.333812714, 0.288573951), 1.00: (0.241716921, 0.196478188, 0.151239425)}
    for alpha, w in want.items():
        got = r.u_ltc_pixel((1, 0.5, 1.5), (0.2, 0.1, 0), (0, 0, 1), (0.8, 0.6, 0.4), alpha)
        np.testing.assert_allclose(got, w, rtol=1e-4)
