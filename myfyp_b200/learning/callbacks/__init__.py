from myfyp_b200.learning.callbacks.callback import CallbackFactory, P2PFLCallback
from myfyp_b200.learning.callbacks.fedprox_callback import FedProxCallback
from myfyp_b200.learning.callbacks.scaffold_callback import SCAFFOLDCallback

CallbackFactory.register_callback("pytorch", SCAFFOLDCallback)
CallbackFactory.register_callback("pytorch", FedProxCallback)

__all__ = ["CallbackFactory", "P2PFLCallback", "SCAFFOLDCallback", "FedProxCallback"]
