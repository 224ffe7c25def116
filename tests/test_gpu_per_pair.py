"""Per-(timestep, layer) heatmaps — BASELINE config 4's mode ("per-layer and per-timestep heatmap
accumulation").  The reference produces them as `heatmaps[timesteps][:, layer_indices]` before
the mean (concept_attention_pipeline.py:76-82; experiments/per_layer_segmentation); here the
same K3/K4 launch that updates the running sum also writes the pair's own slot of one
[Nt, L, 2, B, C, P] fp32 buffer, addressed through a device-side slot index so that a captured
forward is replayed for every timestep."""
import numpy as np
import pytest
import torch

import golden_inputs as gi
import oracle
from ref_runner import PAIRS, rel_map

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


def _engine(cfg):
    from conceptattention_b200.dit import FluxDiTEngine
    from conceptattention_b200.weights import CONFIGS, init_state_dict
    p = CONFIGS[cfg]
    sd = init_state_dict(p, seed=gi.WEIGHT_SEED, norm_scale_jitter=0.25)
    return p, sd, FluxDiTEngine(p, sd, device="cuda:0")


@pytest.mark.parametrize("norm", ["softmax", "sparsemax", "entmax15"])
def test_per_pair_maps_match_oracle_per_timestep_and_layer(norm):
    """tiny-dev (guidance-distilled architecture like config 4), 3 steps x 2 layers: every slot vs
    oracle.heatmaps_from_vectors on that single (t, l) applied to the engine's own vectors
    (north_star tolerance 1e-2; measured ~1e-6), and mean over slots == the running accumulator."""
    from conceptattention_b200.dit import HeatmapAccumulator
    p, sd, eng = _engine("tiny-dev")
    inp = gi.dit_inputs(p, gi.TINY_SIDE, gi.TINY_T, gi.TINY_C, gi.INPUT_SEED, concept_scale=3.0)
    nsteps, layers = 3, [0, 1]
    ts = oracle.get_schedule(nsteps, inp["img"].shape[1], shift=True)
    P, C = inp["img"].shape[1], gi.TINY_C
    acc = HeatmapAccumulator(1, C, P, "cuda:0", layers, list(range(nsteps)), softmax=(norm == "softmax"),
                             attention_norm=norm, per_pair=True)
    img = inp["img"].cuda().to(BF16)
    g = torch.full((1,), 3.5, device="cuda", dtype=BF16)
    collected = {k: [] for k in oracle.VECTOR_KEYS}
    for step, (tc, tp) in enumerate(zip(ts[:-1], ts[1:])):
        acc.begin_step(step)
        tv = torch.full((1,), tc, device="cuda", dtype=BF16)
        pred, vecs = eng(img=img, img_ids=inp["img_ids"], txt=inp["txt"], txt_ids=inp["txt_ids"],
                         concepts=inp["concepts"], concept_ids=inp["concept_ids"],
                         concept_vec=inp["concept_vec"], y=inp["vec"], timesteps=tv, guidance=g,
                         accumulator=acc, collect_vectors=True)
        for k in oracle.VECTOR_KEYS:
            collected[k].append(vecs[k].float().cpu())
        img = img + (tp - tc) * pred
    stacked = {k: torch.stack(v, 0) for k, v in collected.items()}
    pp = acc.per_pair_heatmaps()
    mean = acc.finalize()
    assert pp[0].shape == (nsteps, len(layers), 1, C, 12, 12)
    worst = 0.0
    for space, (a, b) in enumerate(PAIRS):
        for t in range(nsteps):
            for j, l in enumerate(layers):
                want = oracle.heatmaps_from_vectors(stacked[a], stacked[b], [l], [t], softmax=(norm == "softmax"),
                                                    attention_norm=norm)
                r = rel_map(pp[space][t, j], want)
                worst = max(worst, r)
                assert r < 1e-2, (norm, space, t, l, r)
                assert (pp[space][t, j].cpu().argmax(1) == want.argmax(1)).float().mean().item() >= 0.999
        torch.testing.assert_close(pp[space].mean(dim=(0, 1)), mean[space], rtol=1e-5, atol=1e-6)
    print(f"per-pair maps [{norm}]: worst max-rel over {2 * nsteps * len(layers)} slots {worst:.3e}")


def test_per_pair_through_the_public_api_and_graph_replay():
    """generate_image(return_per_pair=True): slots identical launch by launch and under CUDA-graph
    replay (one captured forward for all timesteps, slot index read on the device), timestep /
    layer subsets land in the right slots, and the per-pair buffer does not disturb the mean."""
    from conceptattention_b200 import ConceptAttentionFluxPipeline
    pipe = ConceptAttentionFluxPipeline(model_name="tiny-dev", device="cuda:0")
    eng = pipe.flux_generator.model
    kw = dict(width=256, height=256, num_inference_steps=4, guidance=3.5, return_pil_heatmaps=False,
              decode_image=False)
    concepts = ["dragon", "rock", "sky"]

    def run(seed, **extra):
        o = pipe.generate_image("a dragon on a rock", concepts, seed=seed, **kw, **extra)
        return (o.concept_heatmaps, o.cross_attention_maps,
                None if o.per_pair_concept_heatmaps is None else o.per_pair_concept_heatmaps.cpu().numpy(),
                None if o.per_pair_cross_attention_maps is None else o.per_pair_cross_attention_maps.cpu().numpy())

    eng.use_graphs = False
    plain = run(0, layer_indices=[0, 1])
    ref = [run(s, layer_indices=[0, 1], return_per_pair=True) for s in (0, 1)]
    sub = run(2, layer_indices=[1], timesteps=[1, 3], return_per_pair=True)
    assert ref[0][2].shape == (4, 2, 1, 3, 16, 16) and sub[2].shape == (2, 1, 1, 3, 16, 16)
    np.testing.assert_array_equal(plain[0], ref[0][0])               # the mean is untouched
    np.testing.assert_allclose(ref[0][2].mean(axis=(0, 1))[0], ref[0][0], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(sub[3].mean(axis=(0, 1))[0], sub[1], rtol=1e-5, atol=1e-6)
    assert not np.array_equal(ref[0][2][0], ref[0][2][3])            # timesteps really differ
    eng.use_graphs = True
    eng._graphs.clear(); eng._graph_seen.clear()
    got = [run(s, layer_indices=[0, 1], return_per_pair=True) for s in (0, 1)]
    assert len(eng._graphs) == 1, len(eng._graphs)                   # ONE graph serves the 4 timesteps
    got_sub = run(2, layer_indices=[1], timesteps=[1, 3], return_per_pair=True)
    for a, b in zip(got + [got_sub], ref + [sub]):
        for x, y in zip(a, b):
            np.testing.assert_allclose(x, y, rtol=1e-5, atol=1e-7)
