"""Import stub for ``pytorch_transformers`` (absent on this image).  The CV baseline only
needs the import in the reference's data_utils/fed_persona.py:15 to succeed."""
import _ref_compat  # noqa: F401

CONFIG_NAME = "config.json"
WEIGHTS_NAME = "pytorch_model.bin"


def cached_path(url_or_filename, cache_dir=None):
    raise RuntimeError("no network: cannot fetch %s" % (url_or_filename,))


class _Unavailable:
    def __init__(self, *a, **kw):
        raise RuntimeError("pytorch_transformers is not installed; GPT-2 baseline uses the adapter in run_ref.py")

    @classmethod
    def from_pretrained(cls, *a, **kw):
        raise RuntimeError("pytorch_transformers is not installed")


AdamW = OpenAIGPTDoubleHeadsModel = OpenAIGPTTokenizer = GPT2DoubleHeadsModel = GPT2Tokenizer = _Unavailable
