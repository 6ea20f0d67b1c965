"""s2p_b200 -- B200-native (sm_100a) implementation of the S2P state-conditioned
generator hot path (dsshim0125/s2p, --netG=s2p).  See DESIGN.md."""
from .options import make_options  # noqa: F401
from .networks.generator import S2PGeneratorB200, define_G  # noqa: F401
from .networks.discriminator import MultiscaleDiscriminatorB200, define_D  # noqa: F401
from .networks.vgg import VGG19FeaturesB200  # noqa: F401

__all__ = ["make_options", "define_G", "define_D", "S2PGeneratorB200", "MultiscaleDiscriminatorB200", "VGG19FeaturesB200"]
