"""The four L0 function interfaces of the reference (src/ActivationFunction.java:14,
src/ActivationPrime.java:14, src/LossFunction.java:16, src/LossFunctionPrime.java:13) and the named
singletons that map to kernel enums.  A singleton is still a plain scalar callable (so user code
that calls `.func(z)` keeps working) but also carries the kernel id the backend needs; arbitrary
lambdas cannot run on the GPU and are rejected by the net constructors (no CPU fallback)."""
import math


class ActivationFunction:
    """double -> Double (src/ActivationFunction.java:14)."""
    kernel_id = None
    param = 0.0

    def func(self, a):
        raise NotImplementedError


class ActivationPrime:
    """derivative, evaluated on the PRE-activation (S:277, G:270) (src/ActivationPrime.java:14)."""
    kernel_id = None
    param = 0.0

    def func(self, a):
        raise NotImplementedError


class LossFunction:
    """loss(calculated, expected) (src/LossFunction.java:16)."""
    kernel_id = None

    def loss(self, calculated, expected):
        raise NotImplementedError


class LossFunctionPrime:
    """func(calculated, expected) (src/LossFunctionPrime.java:13)."""
    kernel_id = None

    def func(self, calculated, expected):
        raise NotImplementedError


class _Act(ActivationFunction):
    def __init__(self, kid, f, param=0.0, name=""):
        self.kernel_id, self._f, self.param, self.name = kid, f, param, name

    def func(self, a):
        return self._f(a)

    __call__ = func

    def __repr__(self):
        return "ActivationFunction.%s" % self.name


class _ActPrime(ActivationPrime):
    def __init__(self, kid, f, param=0.0, name=""):
        self.kernel_id, self._f, self.param, self.name = kid, f, param, name

    def func(self, a):
        return self._f(a)

    __call__ = func

    def __repr__(self):
        return "ActivationPrime.%s" % self.name


def _sigmoid(z):
    return 1.0 / (1.0 + math.exp(-z))


def leaky_relu(alpha=0.01):
    """the in-tree pair, src/MNISTTrainer.java:234-235"""
    return (_Act(0, lambda a: a if a > 0 else alpha * a, alpha, "LEAKY_RELU"),
            _ActPrime(0, lambda a: alpha if a <= 0.0 else 1.0, alpha, "LEAKY_RELU_PRIME"))


LEAKY_RELU, LEAKY_RELU_PRIME = leaky_relu(0.01)
SIGMOID = _Act(1, _sigmoid, 0.0, "SIGMOID")  # doc/backprop.pdf
SIGMOID_PRIME = _ActPrime(1, lambda z: _sigmoid(z) * (1.0 - _sigmoid(z)), 0.0, "SIGMOID_PRIME")
RELU = _Act(2, lambda a: a if a > 0 else 0.0, 0.0, "RELU")
RELU_PRIME = _ActPrime(2, lambda a: 0.0 if a <= 0.0 else 1.0, 0.0, "RELU_PRIME")
TANH = _Act(3, math.tanh, 0.0, "TANH")
TANH_PRIME = _ActPrime(3, lambda z: 1.0 - math.tanh(z) ** 2, 0.0, "TANH_PRIME")
IDENTITY = _Act(4, lambda a: a, 0.0, "IDENTITY")
IDENTITY_PRIME = _ActPrime(4, lambda a: 1.0, 0.0, "IDENTITY_PRIME")


class _HalfSquaredError(LossFunction):
    kernel_id = 0

    def loss(self, calculated, expected):
        return 0.5 * (calculated - expected) ** 2


class _HalfSquaredErrorPrime(LossFunctionPrime):
    kernel_id = 0

    def func(self, calculated, expected):
        return calculated - expected


HALF_SQUARED_ERROR = _HalfSquaredError()  # doc/backprop.pdf: L = sum 1/2 (y - yhat)^2
HALF_SQUARED_ERROR_PRIME = _HalfSquaredErrorPrime()


def activation_pair_id(func, prime, what):
    """kernel id + parameter of a built-in pair, or ValueError (IllegalArgumentException in Java)."""
    kid = getattr(func, "kernel_id", None)
    pid = getattr(prime, "kernel_id", None)
    if kid is None or pid is None:
        raise ValueError("%s activation must be one of the built-in singletons (functions.LEAKY_RELU, SIGMOID, ...): "
                         "arbitrary lambdas cannot run in the GPU backend and there is no CPU fallback" % what)
    if kid != pid or getattr(func, "param", 0.0) != getattr(prime, "param", 0.0):
        raise ValueError("%s activation and derivative do not belong together" % what)
    return kid, float(getattr(func, "param", 0.0))


def loss_pair_id(loss, prime):
    kid = getattr(loss, "kernel_id", None)
    pid = getattr(prime, "kernel_id", None)
    if kid is None or pid is None or kid != pid:
        raise ValueError("loss must be one of the built-in singletons (functions.HALF_SQUARED_ERROR): "
                         "arbitrary lambdas cannot run in the GPU backend and there is no CPU fallback")
    return kid
