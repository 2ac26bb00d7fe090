"""Hyphenated script name of the reference (trainscripts/imagesliders/train_lora-scale.py); see train_lora_scale.py."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sliders_conceptmod_b200.train_lora_scale import build_parser, main  # noqa: E402

if __name__ == "__main__":
    main(build_parser().parse_args())
