"""Generate the reference-format index fixtures (tests/golden/ref_*.idx): files exactly as the REFERENCE's `speq index` writes
them (src/fm_indexer.cpp:39-50) -- a cereal::BinaryOutputArchive of the metadata and of the real vendored sdsl-lite csa_wt /
sd_vector objects, which serialise themselves.

Run in the build container, where /root/reference exists:   python tests/golden/make_ref_idx.py
oracle/_ref/speq_ref `write-idx` does the writing (it links the real sdsl + cereal headers; seqan3::fm_index itself does not
compile under g++ 13, so the member order of its serialize function, fm_index.hpp:486-514, is restated there).  The fixtures
pin the product's reader (speqg_refidx_*, speq_b200/csrc/refidx.cpp) to bytes the reference's own types wrote.
"""
from __future__ import annotations

import subprocess
import sys
import tempfile
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import helpers  # noqa: E402

FIXTURES = {"ref_global_single.idx": "global_single", "ref_refs_with_n.idx": "refs_with_n"}


def main() -> None:
    ref = helpers.ref_binary()
    assert ref is not None, "needs /root/reference to build oracle/_ref/speq_ref"
    for out, name in FIXTURES.items():
        fx = helpers.load_golden(name)
        with tempfile.TemporaryDirectory() as d:
            refs, groups, _, _ = helpers.golden_files(Path(d), fx)
            subprocess.run([str(ref), "write-idx", "refs.fa", "groups.txt", str(helpers.GOLDEN / out)], check=True, cwd=d)
        print(out, (helpers.GOLDEN / out).stat().st_size, "bytes")


if __name__ == "__main__":
    main()
