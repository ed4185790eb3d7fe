"""Spatial light distribution with many lights: the per-voxel tables in their sparse form (rows only for
the voxels that can hold a surface point) give the film the full table gives, bit for bit; the 1152-light
fixture (tests/golden: manylights, an emissive mesh) is rendered through it against the reference."""
import numpy as np
import pytest

from conftest import blob_path, golden

pytestmark = pytest.mark.gpu


def _film(capi, name, sparse):
    g = golden(name)
    ctx = capi.Context(0)
    ctx.set_option("sparse_voxels", int(sparse))
    sf = capi.SceneFile(blob_path(name))
    ctx.upload(sf)
    sf.close()
    w, h = [int(v) for v in g["res"]]
    ctx.render(0, int(g["spp"]))
    planes = ctx.film_pack_dat(w, h)
    rows = ctx.light_table_rows()
    ctx.close()
    return planes, rows


@pytest.mark.parametrize("name", ["cornell64", "texbox", "uber_box", "envbox_const"])
def test_sparse_tables_give_the_same_film(capi, name):
    dense, n_dense = _film(capi, name, False)
    sparse, n_sparse = _film(capi, name, True)
    assert 0 < n_sparse < n_dense, (n_sparse, n_dense)
    assert np.array_equal(dense, sparse)


def test_many_lights_switch_to_sparse_tables(capi):
    planes, rows = _film(capi, "manylights", False)   # not forced: 1152 lights x 64^3 voxels pass the dense limit
    g = golden("manylights")
    assert 0 < rows < 64 ** 3 // 4
    ref = g["dat"].astype(np.float64)
    tot_ref, tot_mine = ref.sum(axis=0), planes.sum(axis=0)
    flipped = np.abs(tot_mine - tot_ref) > 1e-3 * np.maximum(tot_ref, ref.mean() * ref.shape[0])
    assert flipped.sum() <= max(1, int(5e-4 * flipped.size)), f"{flipped.sum()} pixels differ"
    keep = ~flipped
    rel_rmse = np.sqrt((((planes - ref) ** 2)[:, keep]).mean(axis=1)) / np.maximum(ref[:, keep].mean(axis=1), 1e-12)
    assert rel_rmse.max() < 0.005
