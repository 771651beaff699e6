"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/pgen.h declares, and its
argument/IO validation fails loudly without a GPU (no compute calls here)."""
import os
import re

import pytest

from textsummarization_on_flink_b200 import native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "pgen.h")).read()
    return sorted(set(re.findall(r"PGEN_API[^;]*?\b(pgen_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = native.lib()
    names = _declared()
    assert len(names) >= 18
    for n in names:
        assert hasattr(L, n), n
    assert sorted(native.EXPORTS) == names


def test_create_rejects_bad_flags_without_touching_the_gpu(tmp_path):
    for flags, code in [("p --mode=train", 5), ("p --mode=banana", 1), ("p --mode=decode --beam_size=x", 1),
                        ("p --mode=decode --vocab_path=/nonexistent/vocab", 2)]:
        with pytest.raises(native.PgenError) as e:
            native.Decoder(flags)
        assert e.value.code == code, flags
    v = tmp_path / "vocab"
    v.write_text("a 1\na 2\n")
    with pytest.raises(native.PgenError) as e:
        native.Decoder("p --mode=decode --vocab_path=%s" % v)
    assert e.value.code == 2 and "Duplicated word" in str(e.value)


def test_no_cpu_fallback(tmp_path):
    """Without a CUDA device pgen_create must fail (PGEN_ERR_CUDA), never decode on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    v = tmp_path / "vocab"
    v.write_text("".join("w%d 1\n" % i for i in range(50)))
    with pytest.raises(native.PgenError) as e:
        native.Decoder("p --mode=decode --vocab_path=%s --log_root=%s --exp_name=x" % (v, tmp_path))
    assert e.value.code == 3 and "no CPU fallback" in str(e.value)


def test_device_count_without_a_gpu():
    """pgen_device_count() is what PgenNative.deviceCount() binds: 0 when the process sees no CUDA device."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    assert native.device_count() == 0
