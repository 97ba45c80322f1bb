"""The oracle (oracle/rt_oracle.cpp) is pinned against the reference itself:
* committed fixtures produced by the unmodified reference build (tests/golden/make_golden.py) — BIT-EXACT
  float64 framebuffers, ids, hit distances and per-ray records;
* where oracle/_ref/ref_driver is present, a live comparison on scene files."""
import numpy as np
import pytest

from conftest import golden_names, golden_oracle, scene_path


@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_reference_fixture_standard(oa, goldens, name):
    g = goldens[name]
    got = golden_oracle(oa, g).render("standard", want_ids=True)
    assert got["fb"].shape == g["std_fb"].shape
    assert np.array_equal(got["fb"], g["std_fb"]), "float64 framebuffer differs from the reference's"
    assert np.array_equal(got["ids"], g["std_ids"])
    assert np.array_equal(got["t"], g["std_t"])


@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_reference_fixture_paper(oa, goldens, name):
    g = goldens[name]
    got = golden_oracle(oa, g).render("paper", want_ids=True)
    assert np.array_equal(got["fb"], g["paper_fb"])
    assert np.array_equal(got["ids"], g["paper_ids"])
    assert np.array_equal(got["t"], g["paper_t"])


@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_reference_fixture_rays(oa, goldens, name):
    g = goldens[name]
    got = golden_oracle(oa, g).trace_rays(g["rays"])
    assert np.array_equal(got[:, :14], g["ray_records"][:, :14])
    assert g["ray_records"][:, 0].sum() > 20, "fixture should contain hits"


def test_oracle_band_equals_full_frame(oa, goldens):
    """Rows [y0,y1) rendered after discard() reproduce the full-frame stream (the RNG is consumed in pixel order)."""
    g = goldens["quirks"]
    o = golden_oracle(oa, g)
    H = g["std_fb"].shape[0]
    band = o.render("standard", rows=(10, 20), threads=1)["fb"]
    assert np.array_equal(band[H - 20:H - 10], g["std_fb"][H - 20:H - 10])   # loop row y is stored in row H-1-y
    assert not band[: H - 20].any() and not band[H - 10:].any()


# number_syntax: exponents, -0, 30-digit mantissas, escapes, duplicate keys — the oracle port shares the product's JSON reader
# (csrc/json_min.hpp), ref_driver parses with the reference's nlohmann/json: equal framebuffers mean equal numbers
@pytest.mark.parametrize("scene,mode", [("pokeballs_small", "standard"), ("quirks", "paper"), ("synthetic_b_small", "standard"),
                                        ("number_syntax", "standard"), ("number_syntax", "paper")])
def test_oracle_matches_live_reference(oa, scene, mode):
    if not oa.have_ref():
        pytest.skip("oracle/_ref/ref_driver not built on this machine")
    p = scene_path(scene)
    ref = oa.ref_render(p, mode, want_ids=True)
    got = oa.Oracle.from_file(p).render(mode, want_ids=True)
    assert np.array_equal(got["fb"], ref["fb"])
    assert np.array_equal(got["ids"], ref["ids"])


def test_oracle_directional_lights_match_live_reference(oa):
    """DirectionalLight (scene.h:10-19, shading.cpp:46-76) cannot be written in the JSON format; ref_driver --dirlight pushes
    them into scene.dir_lights of the unmodified reference, the oracle takes them through orc_add_dir_light."""
    if not oa.have_ref():
        pytest.skip("oracle/_ref/ref_driver not built on this machine")
    p = scene_path("pokeballs_small")
    dl = [((0.2, -1.0, -0.4), (0.9, 0.8, 0.7)), ((0.0, 0.0, -1.0), (0.2, 0.2, 0.2)), ((0.0, 0.0, 0.0), (0.5, 0.1, 0.1))]   # the last: zero dir -> +Y fallback
    for mode in ("standard", "paper"):
        ref = oa.ref_render(p, mode, want_ids=True, dir_lights=dl)
        o = oa.Oracle.from_file(p)
        for d, r in dl:
            o.add_dir_light(d, r)
        got = o.render(mode, want_ids=True)
        assert np.array_equal(got["fb"], ref["fb"]) and np.array_equal(got["ids"], ref["ids"])
        plain = oa.Oracle.from_file(p).render(mode)
        assert not np.array_equal(plain["fb"], got["fb"]), "the lights must be visible"
