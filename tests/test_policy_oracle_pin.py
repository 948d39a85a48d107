"""Pin of oracle/policy_oracle.py on the reference's OWN class where one exists.

tests/golden/policy_features.npz was produced by the unmodified LSTMFeaturesExtractor (learned_controllers/networks/
lstm_policy.py:13-97; its stable-baselines3 base class is the 20-line stand-in of oracle/sb3_stub.py) and by
LSTMPolicy / SimpleMLPPolicy.get_policy_kwargs (:107-164): state_dict, observations, features.  The restated extractor must
reproduce the features from that state_dict, and the actor sizes the oracle hard-codes must be the ones the reference
hands to sb3-contrib.  What stays unpinned is the third-party half (sb3-contrib's stateful actor LSTM, pi-MLP wiring,
action head and clipping), restated from its published behaviour."""
import os

import numpy as np
import torch

from conftest import GOLDEN
from oracle import policy_oracle as po


def _load():
    g = np.load(os.path.join(GOLDEN, "policy_features.npz"))
    sd = {k[3:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd/")}
    gen = torch.Generator().manual_seed(99)
    for layer, width in (("l0", 256), ("l1", 256)):      # never read: the extractor's LSTM runs from h0 = 0 on every call
        sd[f"lstm.weight_hh_{layer}"] = torch.randn((1024, width), generator=gen)
    return g, sd


def test_features_extractor_reproduces_the_reference_class():
    g, sd = _load()
    fe = po.LSTMFeaturesExtractorOracle()
    missing = fe.load_state_dict(sd, strict=True)            # same layer names as the reference: loads as it is
    assert not missing.missing_keys and not missing.unexpected_keys
    fe.eval()
    # the sequence branch ([batch, seq, obs] -> last step, :88-91), where weight_hh matters: a small instance of the same class
    small = po.LSTMFeaturesExtractorOracle(features_dim=16, lstm_hidden_size=32, n_lstm_layers=2)
    small.load_state_dict({k[6:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("small/")}, strict=True)
    small.eval()
    with torch.no_grad():
        feats = fe(torch.from_numpy(g["obs"]))
        feats_seq = small(torch.from_numpy(g["obs_seq"]))
        assert float((small(torch.from_numpy(g["obs"][:64])) - torch.from_numpy(g["features_small"])).abs().max()) <= 1e-6
    assert feats.shape == (512, 128) and int(g["features_dim"]) == fe.features_dim == po.FEATURES_DIM
    # same torch ops in the same order: fp32 round-off at most (bit-identical under the generating torch version)
    assert float((feats - torch.from_numpy(g["features"])).abs().max()) <= 1e-6
    assert float((feats_seq - torch.from_numpy(g["features_seq"])).abs().max()) <= 1e-6
    assert float(torch.from_numpy(g["features"]).abs().max()) > 0.05          # not a dead network


def test_bf16_emulation_tracks_the_reference_features():
    """forward_bf16_emulated (what the tcgen05 kernel computes) against the reference's features through the extractor part."""
    g, sd = _load()
    net = po.make_policy("lstm", seed=0)
    net.features_extractor.load_state_dict(sd)
    obs = torch.from_numpy(g["obs"])
    with torch.no_grad():
        act, _ = net(obs, net.initial_state(512), None)
        act_e, _ = po.forward_bf16_emulated(net, obs, net.initial_state(512), None)
    assert float((net.features_extractor(obs) - torch.from_numpy(g["features"])).abs().max()) <= 1e-6
    assert float((act - act_e).abs().max()) <= 5e-3


def test_actor_sizes_are_what_the_reference_hands_to_sb3():
    g, _ = _load()
    assert g["lstm_kwargs"].tolist() == [po.FEATURES_DIM, po.LSTM_HIDDEN, 2]
    assert g["lstm_pi"].tolist() == [128, 64] and g["lstm_vf"].tolist() == [128, 64]
    assert str(g["lstm_activation"]) == "ReLU" and str(g["mlp_activation"]) == "ReLU"
    assert g["mlp_pi"].tolist() == [256, 128, 64]
    lstm, mlp = po.make_policy("lstm"), po.make_policy("mlp")
    pn = lstm.mlp_extractor.policy_net
    assert [pn[0].in_features, pn[0].out_features, pn[2].out_features] == [po.LSTM_HIDDEN, 128, 64]
    assert lstm.lstm_actor.input_size == po.FEATURES_DIM and lstm.lstm_actor.hidden_size == 256 and lstm.lstm_actor.num_layers == 1
    pm = mlp.mlp_extractor.policy_net
    assert [pm[0].in_features, pm[0].out_features, pm[2].out_features, pm[4].out_features] == [18, 256, 128, 64]
    assert all(isinstance(m, torch.nn.ReLU) for m in (pn[1], pn[3], pm[1], pm[3], pm[5]))
