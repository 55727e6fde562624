#!/usr/bin/env python
"""`python methods.py cont_fn style_fn [--epochs ... --gatys ...]` -- the reference's command line
(/root/reference/methods.py:243-271) on the B200 step.  See audio_style_transfer_b200/methods.py."""
from audio_style_transfer_b200.methods import main

if __name__ == '__main__':
    main()
