"""GPU test of W-tiled inference (tiled.py): the eval-mode encode / decode of a long
utterance computed on column tiles with halos must equal the untiled call BIT FOR BIT (same taps, same kernels, same
order), on the production engine at the config-5 length and on the fp32 engine."""
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import neural_sound_generation_b200 as nsg
    from neural_sound_generation_b200 import tiled as T
    from oracle import vqvae_ref as R


@pytest.mark.parametrize('precision,frames', [('auto', 861), ('auto', 256), ('fp32', 203)])
def test_tiled_encode_decode_equal_untiled(precision, frames):
    p, b = R.make_state(1, 64, 512, 3, codebook='spread')
    m = nsg.VQVAE(1, 64, 512)
    sd = dict(p)
    sd.update(b)
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    m = m.to('cuda').eval()
    if precision == 'fp32':
        m.set_precision('fp32')
    x = R.make_mel(1, 80, frames, 9).unsqueeze(1).to('cuda')
    with torch.no_grad():
        want_lat = m.encode(x)
        want_xt = m.decode(want_lat)
    for n in (2, 3, 5):
        lat, bounds = T.encode_tiled(m, x, n)
        assert len(bounds) == n and lat.shape == want_lat.shape
        assert torch.equal(lat, want_lat), (precision, frames, n, (lat != want_lat).sum().item())
        xt, _ = T.decode_tiled(m, want_lat, n)
        assert xt.shape == want_xt.shape
        assert torch.equal(xt, want_xt), (precision, frames, n, (xt - want_xt).abs().max().item())
    lat, xt = T.encode_decode_sharded(m, x, n_tiles=4)           # no process group: every tile on this device
    assert torch.equal(lat, want_lat) and torch.equal(xt, want_xt)
