"""Generate golden vectors by running the REFERENCE's own Python modules.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
Outputs tests/golden/*.npz.  The reference cannot travel to the GPU box, the
fixtures do.  Inputs are not stored: they are re-derived from seeds by
tests/golden_inputs.py (CPU torch RNG is reproducible for a given torch version;
the generating torch version is recorded in each file).

Four third-party modules the reference imports eagerly are absent from this
image and are stubbed (none is on the hot path): entmax, matplotlib, imwatermark,
fire.  The reference source files are imported unmodified.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def install_stubs():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    def _absent(*a, **k):
        raise RuntimeError("entmax is not installed (parity unpinned)")
    mod("entmax", entmax15=_absent, sparsemax=_absent)
    mpl = mod("matplotlib")
    mpl.pyplot = mod("matplotlib.pyplot")
    mpl.animation = mod("matplotlib.animation")

    class WatermarkEncoder:
        def set_watermark(self, *a, **k): pass
        def encode(self, x, *a, **k): return x
    mod("imwatermark", WatermarkEncoder=WatermarkEncoder)
    mod("fire", Fire=lambda *a, **k: None)


def main():
    install_stubs()
    sys.path.insert(0, "/root/reference")
    import concept_attention  # noqa: F401  (runs the reference's eager imports)
    from concept_attention.modified_flux_dit import ModifiedFluxDiT, FluxParams as RefParams
    from concept_attention.concept_attention_pipeline import compute_heatmaps_from_vectors
    from concept_attention.flux.src.flux.sampling import denoise, get_schedule, get_noise, prepare
    from concept_attention.utils import linear_normalization
    from concept_attention.segmentation import SegmentationAbstractClass, add_noise_to_image

    import golden_inputs as gi
    from conceptattention_b200.weights import CONFIGS, init_state_dict

    torch.set_grad_enabled(False)
    meta = dict(torch_version=torch.__version__, reference_commit="401dc36")

    # ---- 1. DiT forward: vectors of every double block + prediction ----------
    for cfg_name in ("tiny", "tiny-dev"):
        p = CONFIGS[cfg_name]
        sd = init_state_dict(p, seed=gi.WEIGHT_SEED, norm_scale_jitter=0.25)
        model = ModifiedFluxDiT(RefParams(**p.asdict()))
        missing, unexpected = model.load_state_dict(sd, strict=True)
        assert not missing and not unexpected
        model.eval()
        inp = gi.dit_inputs(p, side=gi.TINY_SIDE, T=gi.TINY_T, C=gi.TINY_C, seed=gi.INPUT_SEED)
        t_vec = torch.full((1,), 0.75)
        g_vec = torch.full((1,), 3.5)
        pred, vecs = model(img=inp["img"], img_ids=inp["img_ids"], txt=inp["txt"],
                           txt_ids=inp["txt_ids"], concepts=inp["concepts"],
                           concept_ids=inp["concept_ids"], concept_vec=inp["concept_vec"],
                           y=inp["vec"], timesteps=t_vec, guidance=g_vec)
        out = {k: v.numpy() for k, v in vecs.items()}
        out["pred"] = pred.numpy()
        # ablation branches of the second attention (block-level flags)
        for name, kw in gi.ABLATIONS.items():
            _, v2 = model(img=inp["img"], img_ids=inp["img_ids"], txt=inp["txt"],
                          txt_ids=inp["txt_ids"], concepts=inp["concepts"],
                          concept_ids=inp["concept_ids"], concept_vec=inp["concept_vec"],
                          y=inp["vec"], timesteps=t_vec, guidance=g_vec,
                          stop_after_multimodal_attentions=True, joint_attention_kwargs=kw)
            out["ablation_" + name] = v2["output_space_concept_vectors"].numpy()
        np.savez_compressed(os.path.join(HERE, f"dit_{cfg_name}.npz"), **out, **meta)
        print(cfg_name, {k: v.shape for k, v in out.items()})

        # ---- 2. full denoise loop (2 steps) through the reference sampler ----
        if cfg_name == "tiny":
            ts = get_schedule(2, inp["img"].shape[1], shift=False)
            x, _, stacked = denoise(model, img=inp["img"], img_ids=inp["img_ids"], txt=inp["txt"],
                                    txt_ids=inp["txt_ids"], vec=inp["vec"], timesteps=ts,
                                    guidance=0.0, concepts=inp["concepts"],
                                    concept_ids=inp["concept_ids"], concept_vec=inp["concept_vec"])
            np.savez_compressed(
                os.path.join(HERE, "denoise_tiny.npz"), img=x.numpy(), schedule=np.array(ts),
                out_img=stacked["output_space_image_vectors"].numpy(),
                out_con=stacked["output_space_concept_vectors"].numpy(),
                cross_img=stacked["cross_attention_image_vectors"].numpy(),
                cross_con=stacked["cross_attention_concept_vectors"].numpy(), **meta)

    # ---- 3. compute_heatmaps_from_vectors, unmodified (needs P == 4096) ------
    hv = gi.heatmap_inputs()
    res = {}
    for tag, (iv, cv) in hv.items():
        res[tag + "_softmax"] = compute_heatmaps_from_vectors(
            iv, cv, layer_indices=gi.HM_LAYERS, timesteps=gi.HM_TIMESTEPS,
            softmax=True).numpy()
        res[tag + "_softmax_normcon"] = compute_heatmaps_from_vectors(
            iv, cv, layer_indices=gi.HM_LAYERS, timesteps=gi.HM_TIMESTEPS, softmax=True,
            normalize_concepts=True).numpy()
    x = torch.randn(3, 5, 7, generator=torch.Generator().manual_seed(5))
    res["linear_normalization"] = linear_normalization(x, dim=-2).numpy()
    np.savez_compressed(os.path.join(HERE, "heatmaps.npz"), **res, **meta)

    # ---- 4. heatmap -> mask (segmentation.py:33-83) and schedule / noise -----
    class Seg(SegmentationAbstractClass):
        def segment_individual_image(self, image, concepts, caption, **kw):
            return image, None
    coeff = gi.mask_inputs()
    masks_plain, _, _ = Seg()([coeff], None, ["a", "b", "c", "d", "e"], [""], apply_blur=False)
    masks_blur, _, _ = Seg()([coeff], None, ["a", "b", "c", "d", "e"], [""], apply_blur=True)
    lat = get_noise(1, 256, 256, device="cpu", dtype=torch.float32, seed=7)
    noised, ts_tail = add_noise_to_image(torch.zeros(1, 16, 32, 32) + 0.5, num_steps=4,
                                         noise_timestep=2, seed=3, width=256, height=256,
                                         device="cpu", is_schnell=False)
    np.savez_compressed(
        os.path.join(HERE, "masks.npz"),
        masks_plain=masks_plain[0].numpy(), masks_blur=masks_blur[0].numpy(),
        argmax=coeff.argmax(dim=0).numpy(),
        schedule_shift=np.array(get_schedule(4, 1024, shift=True)),
        schedule_noshift=np.array(get_schedule(4, 1024, shift=False)),
        noise=lat.numpy(), noised=noised.float().numpy(), noised_tail=np.array(ts_tail), **meta)
    print("golden written to", HERE)


if __name__ == "__main__":
    main()
