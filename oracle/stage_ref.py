"""Stage the reference's own Python for the path into `oracle/_ref/` (git-ignored).  TEST INFRASTRUCTURE.

The reference (qingyundou/WaveRNN) is pure Python: there is nothing to compile, its "build" for the timed CPU
baseline is its source files as they lie under `/root/reference`.  `__graft_entry__.build()` calls `stage()` in
the build container, where `/root/reference` exists; the copies land in `oracle/_ref/reference/` which is
listed in `.gitignore` (never enters history) but not in `.gpurunignore`, so it travels to the GPU box like the
built `.so` does.  There `bench.py --impl reference` and the `cpu_baseline` leg time the reference's un-patched
`WaveRNN.generate` (`models/fatchord_version.py:169-264`) on the host cores (`kind: "reference"`), and the call-
compatibility test drives the reference's own `gen_wavernn.gen_from_file` (`gen_wavernn.py:38-65`) against the
drop-in class.  When the staged tree is absent both fall back to the oracle port (`kind: "port"`).

Nothing under `wavernn_b200/` reads this directory.
"""
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, '_ref', 'reference')

# what `models/fatchord_version.py`, `models/deepmind_version.py` and `gen_wavernn.py` import, plus hparams
FILES = [
    'hparams.py', 'gen_wavernn.py', 'LICENSE.txt',
    'models/__init__.py', 'models/fatchord_version.py', 'models/deepmind_version.py',
    'utils/__init__.py', 'utils/distribution.py', 'utils/dsp.py', 'utils/display.py', 'utils/dataset.py',
    'utils/paths.py', 'utils/files.py',
    # imported by utils/dataset.py (text front end; only its import has to succeed)
    'utils/text/__init__.py', 'utils/text/cleaners.py', 'utils/text/cmudict.py', 'utils/text/numbers.py',
    'utils/text/symbols.py', 'utils/text/recipes.py', 'utils/text/LICENSE',
]


def stage(src='/root/reference') -> bool:
    """Copy the listed files if the reference tree is present.  -> True when `oracle/_ref/reference` is usable."""
    if os.path.isfile(os.path.join(src, 'models', 'fatchord_version.py')):
        for rel in FILES:
            s, d = os.path.join(src, rel), os.path.join(DEST, rel)
            if not os.path.isfile(s):
                continue
            os.makedirs(os.path.dirname(d), exist_ok=True)
            shutil.copyfile(s, d)
    return os.path.isfile(os.path.join(DEST, 'models', 'fatchord_version.py'))


if __name__ == '__main__':
    print('staged' if stage() else 'reference tree not present; nothing staged')
