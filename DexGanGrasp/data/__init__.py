from .. import _overlay

_overlay(__path__, "data")
