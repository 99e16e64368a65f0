"""CPU: host logic of the input pipeline (token block layout) and that it refuses to run without the device."""
import pytest
import torch


def test_token_layout_is_packed_in_key_order_and_16_byte_aligned():
    from blt_vqg_b200.batch import TOKEN_KEYS, token_layout
    shapes = {"questions": (7, 20), "answers": (7, 5), "posteriors": (7, 21), "answer_types_for_input": (7, 3)}
    lay, total = token_layout(shapes)
    assert list(lay) == [k for k in TOKEN_KEYS]
    end = 0
    for k, (off, B, n) in lay.items():
        assert off % 4 == 0 and off >= end and (B, n) == shapes[k]
        end = off + B * n
    assert total % 4 == 0 and total >= end
    lay2, total2 = token_layout({"questions": (7, 20)})
    assert lay2 == {"questions": (0, 7, 20)} and total2 == 140
    # per-sample vectors (answer lengths, category labels) follow the token tensors
    lay3, total3 = token_layout({"alengths": (7,), "questions": (7, 20), "answer_types": (7,)})
    assert list(lay3) == ["questions", "alengths", "answer_types"]
    assert lay3["alengths"] == (140, 7, 0) and lay3["answer_types"] == (148, 7, 0) and total3 == 156


def test_input_pipeline_has_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from blt_vqg_b200 import _lib
    from blt_vqg_b200.batch import BatchStager, FeatureCache
    with pytest.raises(_lib.B200iqError):
        BatchStager()
    with pytest.raises(_lib.B200iqError):
        FeatureCache(8, 64)
