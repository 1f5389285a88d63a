"""mlcomp_b200 - B200 implementation of MLComp's post-inference scoring and
post-processing path.  ``install_as_mlcomp()`` makes the
package importable under the reference's module names."""
from .compat import install_as_mlcomp  # noqa: F401

__all__ = ['install_as_mlcomp']
