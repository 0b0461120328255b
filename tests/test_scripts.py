"""The example launch scripts must only use flags the parser defines (the reference's imagenet.sh
does not: SURVEY.md 2.1), and the package metadata must import."""
import os
import re
import shlex

import pytest

from commefficient_b200.utils import parse_args

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _flags_of(script):
    text = open(os.path.join(ROOT, "scripts", script)).read()
    text = "\n".join(l for l in text.splitlines() if not l.lstrip().startswith("#"))
    text = text.replace("\\\n", " ")
    text = re.sub(r"\$\(\([^)]*\)\)", "8", text)                   # $((100 * NGPU)) -> a number
    text = text.replace('"$DATA"', "/data").replace('"$@"', " ").replace('"${COMMON[@]}"', " ")
    text = text.replace("COMMON=(", " ").replace(")", " ")
    toks = text.split()
    start = toks.index("--dataset_name")
    out = []
    for t in toks[start:]:
        if t in ("if", "fi", "exec", "then", "python"):
            break
        out.append(t)
    return out


@pytest.mark.parametrize("script,gpt2", [("cifar10_fetchsgd.sh", False), ("imagenet.sh", False),
                                         ("persona_gpt2.sh", True)])
def test_script_flags_parse(script, gpt2):
    flags = _flags_of(script)
    assert "--mode" in flags
    args = parse_args(default_lr=0.4, argv=flags)
    assert args.mode in ("sketch", "uncompressed")


def test_setup_metadata_is_importable():
    src = open(os.path.join(ROOT, "setup.py")).read()
    compile(src, "setup.py", "exec")
    assert "build_native" in src and "commefficient_b200" in src
