"""Stand-in for the slice of ``mx.operator`` the reference's custom ops use
(CustomOp / CustomOpProp / register; MXNet itself is not vendored by the reference,
README.md:29).  With MXNet installed the same classes can be registered with the real
``mx.operator.register`` instead -- see INTEGRATION.md."""

_REGISTRY = {}


class CustomOp(object):
    def forward(self, is_train, req, in_data, out_data, aux):
        raise NotImplementedError

    def backward(self, req, out_grad, in_data, out_data, in_grad, aux):
        raise NotImplementedError

    def assign(self, dst, req, src):
        """mx.operator.CustomOp.assign: req in {'null','write','inplace','add'}."""
        if req == "null":
            return
        if req in ("write", "inplace"):
            dst[:] = src
        elif req == "add":
            dst[:] += src
        else:
            raise ValueError("unknown req %r" % (req,))


class CustomOpProp(object):
    def __init__(self, need_top_grad=True):
        self.need_top_grad_ = need_top_grad

    def need_top_grad(self):
        return self.need_top_grad_

    def list_arguments(self):
        return ["data"]

    def list_outputs(self):
        return ["output"]

    def list_auxiliary_states(self):
        return []

    def infer_shape(self, in_shape):
        return in_shape, [in_shape[0]], []

    def declare_backward_dependency(self, out_grad, in_data, out_data):
        return list(out_grad) + list(in_data) + list(out_data)

    def create_operator(self, ctx, shapes, dtypes):
        raise NotImplementedError


def register(reg_name):
    def deco(prop_cls):
        _REGISTRY[reg_name] = prop_cls
        return prop_cls
    return deco


def get_prop(op_type):
    return _REGISTRY[op_type]


def create(op_type, ctx=None, shapes=None, dtypes=None, **kwargs):
    """What ``mx.symbol.Custom(op_type=..., **kwargs)`` ends up doing: kwargs reach the Prop
    as STRINGS (rcnn/rpn/proposal.py:23-24,171), then create_operator builds the op."""
    prop = _REGISTRY[op_type](**{k: str(v) for k, v in kwargs.items()})
    return prop, prop.create_operator(ctx, shapes, dtypes)
