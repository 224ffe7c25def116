"""f3 / f4 tails.
* f4, weights: the reference's own `ModifiedFluxDiT` at FULL flux-schnell / flux-dev shapes (built on
  the meta device: no 24 GB of host RAM) exports exactly the state-dict keys and shapes the engine
  expects (`weights.expected_keys`); on the GPU a state dict exported by a materialised reference
  model drives the engine directly and the two agree.
* f3, input side: a random-init reference `AutoEncoder` passed as `ae=`; `encode_image(PIL image)`
  resizes and encodes on the device exactly like the reference's `segmentation.encode_image`
  (segmentation.py:115-153) followed by the latent path."""
import numpy as np
import pytest
import torch

from oracle import reference_modules as rm

pytestmark = pytest.mark.skipif(not rm.available(), reason="reference modules not placed (oracle/make_ref.py)")


@pytest.mark.parametrize("name", ["flux-schnell", "flux-dev"])
def test_reference_model_exports_the_keys_and_shapes_the_engine_expects(name):
    from conceptattention_b200.weights import CONFIGS, expected_keys
    ns = rm.load()
    p = CONFIGS[name]
    with torch.device("meta"):
        ref = ns.ModifiedFluxDiT(ns.FluxParams(**p.asdict()))
    got = {k: tuple(v.shape) for k, v in ref.state_dict().items()}
    want = {k: tuple(v) for k, v in expected_keys(p).items()}
    assert set(got) == set(want), (sorted(set(got) ^ set(want))[:8])
    assert got == want
    n = sum(int(np.prod(s)) for s in got.values())
    assert 11.8e9 < n < 12.1e9                               # the 12 B parameter model, not a toy


@pytest.mark.gpu
def test_engine_runs_a_state_dict_exported_by_the_reference_model():
    """flux-schnell architecture at full width (3072, 24 heads) with 2 + 2 blocks: the reference
    model is materialised on the GPU (`to_empty` + its own parameter tensors filled in place), its
    `state_dict()` is handed to the engine unchanged, and both run the same inputs."""
    import dataclasses
    import golden_inputs as gi
    from conceptattention_b200.dit import FluxDiTEngine
    from conceptattention_b200.weights import CONFIGS
    from ref_runner import mae, rel_max
    ns = rm.load()
    p = dataclasses.replace(CONFIGS["flux-schnell"], depth=2, depth_single_blocks=2)
    with torch.device("meta"):
        ref = ns.ModifiedFluxDiT(ns.FluxParams(**p.asdict()))
    ref = ref.to_empty(device="cuda").to(torch.bfloat16).eval()
    g = torch.Generator(device="cuda").manual_seed(3)
    with torch.no_grad():
        for k, v in ref.state_dict().items():
            if k.endswith("scale"):
                v.copy_(1 + 0.1 * torch.randn(v.shape, device="cuda", generator=g))
            elif k.endswith("bias"):
                v.copy_(0.02 * torch.randn(v.shape, device="cuda", generator=g))
            else:
                v.copy_(torch.randn(v.shape, device="cuda", generator=g) / v.shape[-1] ** 0.5)
    sd = ref.state_dict()
    eng = FluxDiTEngine(p, sd, device="cuda:0")
    inp = {k: v.cuda().to(torch.bfloat16) for k, v in gi.dit_inputs(p, 256, 64, 3, gi.INPUT_SEED).items()}
    t = torch.full((1,), 0.6, device="cuda", dtype=torch.bfloat16)
    with torch.no_grad():
        pred, vecs = eng(img=inp["img"], img_ids=inp["img_ids"], txt=inp["txt"], txt_ids=inp["txt_ids"],
                         concepts=inp["concepts"], concept_ids=inp["concept_ids"],
                         concept_vec=inp["concept_vec"], y=inp["vec"], timesteps=t, collect_vectors=True)
        rpred, rvecs = ref(img=inp["img"], img_ids=inp["img_ids"], txt=inp["txt"], txt_ids=inp["txt_ids"],
                           concepts=inp["concepts"], concept_ids=inp["concept_ids"],
                           concept_vec=inp["concept_vec"], y=inp["vec"], timesteps=t, guidance=None)
    assert pred.shape == rpred.shape
    assert rel_max(pred.float(), rpred.float()) < 6e-2, rel_max(pred.float(), rpred.float())
    for k in vecs:
        a, b = vecs[k].float(), rvecs[k].float()
        assert a.shape == b.shape, k
        assert mae(a, b) < 2e-2 * b.abs().mean().item() + 1e-3, (k, mae(a, b))


@pytest.mark.gpu
def test_encode_image_from_pil_through_a_reference_autoencoder():
    PIL = pytest.importorskip("PIL.Image")
    from conceptattention_b200 import ConceptAttentionFluxPipeline
    ns = rm.load()
    ae_params = ns.autoencoder.AutoEncoderParams(resolution=256, in_channels=3, ch=128, out_ch=3,
                                                 ch_mult=[1, 2, 4, 4], num_res_blocks=2, z_channels=16,
                                                 scale_factor=0.3611, shift_factor=0.1159)   # flux/util.py:49-59
    torch.manual_seed(0)
    ae = ns.autoencoder.AutoEncoder(ae_params).to("cuda").eval()
    pipe = ConceptAttentionFluxPipeline(model_name="tiny", device="cuda:0", ae=ae)
    depth = pipe.flux_generator.params.depth
    rgb = np.random.default_rng(1).integers(0, 256, (300, 420, 3), dtype=np.uint8)
    image = PIL.fromarray(rgb)
    kw = dict(concepts=["cat", "sky", "grass"], prompt="a cat", width=512, height=512,
              layer_indices=list(range(depth)), num_samples=2, return_pil_heatmaps=False, seed=4)
    torch.manual_seed(5)                                      # the VAE samples its posterior
    out = pipe.encode_image(image, **kw)
    # the reference's preprocessing by hand (segmentation.py:123-146): ToTensor, 2x-1, to device,
    # F.interpolate to (height, width), autoencoder.encode — then the latent path of the same call
    x = torch.from_numpy(rgb).permute(2, 0, 1).float().div(255.0).mul(2.0).sub(1.0)[None].cuda()
    x = torch.nn.functional.interpolate(x, (512, 512))
    torch.manual_seed(5)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        latent = ae.encode(x)
    assert latent.shape == (1, 16, 64, 64)
    want = pipe.encode_image(latent, **kw)
    assert out.concept_heatmaps.shape == (3, 32, 32) and out.cross_attention_maps.shape == (3, 32, 32)
    # same numbers up to the convolution algorithms cuDNN picks per call inside the bf16 VAE: its
    # latents differ in the last bits between two calls, which random tiny weights amplify on the
    # peaky cross-attention maps — so: close on average, same winners
    for got, ref in ((out.concept_heatmaps, want.concept_heatmaps),
                     (out.cross_attention_maps, want.cross_attention_maps)):
        assert np.abs(got - ref).mean() < 2e-3, np.abs(got - ref).mean()
        assert (got.argmax(0) == ref.argmax(0)).mean() > 0.97
    assert np.isfinite(out.concept_heatmaps).all() and out.concept_heatmaps.std() > 0
    pil = pipe.encode_image(image, **{**kw, "return_pil_heatmaps": True})
    assert len(pil.concept_heatmaps) == 3 and pil.concept_heatmaps[0].size == (32, 32)
