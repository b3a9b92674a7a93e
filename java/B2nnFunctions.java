/**
 * Named singletons for the four functional interfaces of the reference (ActivationFunction.java:14,
 * ActivationPrime.java:14, LossFunction.java:16, LossFunctionPrime.java:13).  Each still works as a
 * plain Java function (so user code and the CPU nets can call it) and additionally carries the
 * kernel enum of include/b2nn.h.  The GPU nets accept ONLY these instances: an arbitrary lambda
 * cannot run inside a CUDA kernel and there is no CPU fallback, so it is rejected with
 * IllegalArgumentException.
 *
 * MNISTTrainer.java:234-235 -- the one edit to that file -- becomes
 *     new SoftmaxCrossEntropyNeuralNet(dims, B2nnFunctions.LEAKY_RELU, B2nnFunctions.LEAKY_RELU_PRIME)
 */
final class B2nnFunctions {
	private B2nnFunctions() {
	}

	/** implemented by every built-in function: which b2nn_activation / b2nn_loss_kind it is */
	interface KernelId {
		int kernelId();

		default double kernelParam() {
			return 0.0;
		}
	}

	static final class Activation implements ActivationFunction, KernelId {
		private final int id;
		private final double p;

		Activation(int id, double p) {
			this.id = id;
			this.p = p;
		}

		@Override
		public Double func(double a) {
			switch (id) {
				case 0: return a > 0 ? a : p * a;               // MNISTTrainer.java:234
				case 1: return 1.0 / (1.0 + Math.exp(-a));       // doc/backprop.pdf
				case 2: return a > 0 ? a : 0.0;
				case 3: return Math.tanh(a);
				default: return a;
			}
		}

		@Override
		public int kernelId() {
			return id;
		}

		@Override
		public double kernelParam() {
			return p;
		}
	}

	static final class Prime implements ActivationPrime, KernelId {
		private final int id;
		private final double p;

		Prime(int id, double p) {
			this.id = id;
			this.p = p;
		}

		/** evaluated on the PRE-activation, like the reference (SoftmaxCrossEntropyNeuralNet.java:277) */
		@Override
		public Double func(double a) {
			switch (id) {
				case 0: return a <= 0.0 ? p : 1.0;               // MNISTTrainer.java:235
				case 1: {
					double s = 1.0 / (1.0 + Math.exp(-a));
					return s * (1.0 - s);
				}
				case 2: return a <= 0.0 ? 0.0 : 1.0;
				case 3: {
					double t = Math.tanh(a);
					return 1.0 - t * t;
				}
				default: return 1.0;
			}
		}

		@Override
		public int kernelId() {
			return id;
		}

		@Override
		public double kernelParam() {
			return p;
		}
	}

	static final Activation LEAKY_RELU = new Activation(0, 0.01);
	static final Prime LEAKY_RELU_PRIME = new Prime(0, 0.01);
	static final Activation SIGMOID = new Activation(1, 0.0);
	static final Prime SIGMOID_PRIME = new Prime(1, 0.0);
	static final Activation RELU = new Activation(2, 0.0);
	static final Prime RELU_PRIME = new Prime(2, 0.0);
	static final Activation TANH = new Activation(3, 0.0);
	static final Prime TANH_PRIME = new Prime(3, 0.0);
	static final Activation IDENTITY = new Activation(4, 0.0);
	static final Prime IDENTITY_PRIME = new Prime(4, 0.0);

	static Activation leakyRelu(double alpha) {
		return new Activation(0, alpha);
	}

	static Prime leakyReluPrime(double alpha) {
		return new Prime(0, alpha);
	}

	/** L = 1/2 (calculated - expected)^2, L' = calculated - expected (doc/backprop.pdf) */
	static final class HalfSquaredError implements LossFunction, KernelId {
		@Override
		public Double loss(double calculated, double expected) {
			double d = calculated - expected;
			return 0.5 * d * d;
		}

		@Override
		public int kernelId() {
			return 0;
		}
	}

	static final class HalfSquaredErrorPrime implements LossFunctionPrime, KernelId {
		@Override
		public Double func(double calculated, double expected) {
			return calculated - expected;
		}

		@Override
		public int kernelId() {
			return 0;
		}
	}

	static final HalfSquaredError HALF_SQUARED_ERROR = new HalfSquaredError();
	static final HalfSquaredErrorPrime HALF_SQUARED_ERROR_PRIME = new HalfSquaredErrorPrime();

	/** kernel id of a function pair, or IllegalArgumentException for anything that is not built in */
	static KernelId require(Object f, Object prime, String what) {
		if (!(f instanceof KernelId) || !(prime instanceof KernelId)
				|| ((KernelId) f).kernelId() != ((KernelId) prime).kernelId()
				|| ((KernelId) f).kernelParam() != ((KernelId) prime).kernelParam()) {
			throw new IllegalArgumentException(what + ": the GPU backend runs only the built-in function "
					+ "pairs of B2nnFunctions (matching function and derivative); arbitrary lambdas have no "
					+ "kernel and there is no CPU fallback");
		}
		return (KernelId) f;
	}
}
