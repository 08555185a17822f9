"""Import alias for the hyphenated package directory `show-attend-and-tell_b200/`.

`import sat_b200` resolves sub-modules (attention, decoder, synth, ...) from that directory.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "show-attend-and-tell_b200")
__path__.insert(0, _real)
