"""The tcgen05 / TMA dense feature-block tile against plain PyTorch fp32 on the same bf16 inputs."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,d,k", [(128, 64, 16), (1000, 200, 5), (4096, 1024, 64), (333, 72, 130)])
def test_dense_margins_match_fp32_reference(n, d, k):
    from parameter_server_b200.ops.dense import dense_margins, pad_features
    g = torch.Generator(device="cuda").manual_seed(n + d + k)
    x = torch.randn(n, d, generator=g, device="cuda") * 0.5
    w = torch.randn(k, d, generator=g, device="cuda") * 0.2
    label = torch.where(torch.rand(n, generator=g, device="cuda") > 0.5, 1.0, -1.0)
    xb, wb = pad_features(x), pad_features(w)
    y, loss = dense_margins(xb, wb, label, d=d)
    torch.cuda.synchronize()
    ref = xb[:, :d].float() @ wb[:, :d].float().T           # same rounded operands, fp32 accumulation
    assert torch.allclose(y, ref, rtol=2e-3, atol=2e-3), float((y - ref).abs().max())
    ref_loss = torch.log1p(torch.exp(-(label[:, None] * ref).double())).sum(0)
    assert torch.allclose(loss, ref_loss, rtol=1e-4)


def test_score_dense_models_ranks_the_planted_model_first():
    from parameter_server_b200.models.dense_scoring import score_dense_models
    g = torch.Generator(device="cuda").manual_seed(1)
    n, d = 20000, 96
    x = torch.randn(n, d, generator=g, device="cuda")
    w_true = torch.randn(d, generator=g, device="cuda")
    label = torch.where(x @ w_true + 0.3 * torch.randn(n, generator=g, device="cuda") > 0, 1.0, -1.0)
    models = torch.stack([w_true, 0.1 * w_true, torch.randn(d, generator=g, device="cuda"), torch.zeros(d, device="cuda")])
    s = score_dense_models(x, label, models)
    assert s[0]["auc"] > 0.97 and s[0]["auc"] == pytest.approx(s[1]["auc"], abs=2e-3)   # scaling keeps the ranking
    assert abs(s[2]["auc"] - 0.5) < 0.2 and s[3]["logloss"] == pytest.approx(0.6931, abs=1e-3)
    ref = torch.log1p(torch.exp(-(label * (x.bfloat16().float() @ models[0].bfloat16().float())).double())).mean()
    assert s[0]["logloss"] == pytest.approx(float(ref), rel=1e-3)
