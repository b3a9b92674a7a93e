import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.lang.ref.Cleaner;
import java.util.HashMap;
import java.util.Map;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_INT;

/**
 * Holder of one native b2nn_net handle; the bodies of the six NeuralNet methods
 * (src/NeuralNet.java:16-65) forwarded to libb2nn.so.  SoftmaxCrossEntropyNeuralNet and
 * GeneralNeuralNet (same directory) extend it and keep the reference's constructors.
 */
abstract class B2nnNet implements NeuralNet {
	private static final Cleaner CLEANER = Cleaner.create();

	final int[] layerDims;
	final MemorySegment handle;

	B2nnNet(int[] layerDims, int kind, int innerAct, double innerParam, int lastAct, double lastParam, int loss) {
		if (layerDims == null || layerDims.length < 2) {
			throw new IllegalArgumentException("layerDims must name at least 2 layers");
		}
		this.layerDims = layerDims.clone();
		try (Arena a = Arena.ofConfined()) {
			MemorySegment out = a.allocate(ADDRESS);
			// seed 1: the reference hard-codes new Random(1) (SoftmaxCrossEntropyNeuralNet.java:111)
			B2nn.check((int) B2nn.CREATE.invokeExact(B2nn.pack(a, layerDims), layerDims.length, kind, innerAct,
					innerParam, lastAct, lastParam, loss, B2nn.PREC_F64, 0, 1L, -1, out));
			handle = out.get(ADDRESS, 0);
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
		final MemorySegment h = handle;
		CLEANER.register(this, () -> {
			try {
				int ignored = (int) B2nn.DESTROY.invokeExact(h);
			} catch (Throwable ignored) {
			}
		});
	}

	@Override
	public double[] propagate(double[] input) {
		if (input == null || input.length != layerDims[0]) throw new IllegalArgumentException("bad input");
		try (Arena a = Arena.ofConfined()) {
			MemorySegment y = a.allocate(JAVA_DOUBLE, getOutputDim());
			B2nn.check((int) B2nn.PROPAGATE.invokeExact(handle, B2nn.pack(a, input), 1, y));
			return y.toArray(JAVA_DOUBLE);
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
	}

	@Override
	public double calculateLoss(double[] input, double[] expected) {
		if (input == null || expected == null || input.length != layerDims[0] || expected.length != getOutputDim())
			throw new IllegalArgumentException("bad input / expected");
		try (Arena a = Arena.ofConfined()) {
			MemorySegment l = a.allocate(JAVA_DOUBLE);
			B2nn.check((int) B2nn.LOSS.invokeExact(handle, B2nn.pack(a, input), B2nn.pack(a, expected), 1, l));
			return l.get(JAVA_DOUBLE, 0);
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
	}

	/** layout: key i <-> weights between layer i and i+1, [layer i neuron][layer i+1 neuron] (NeuralNet.java:30-39) */
	@Override
	public Map<Integer, double[][]> calculateWeightGradient(double[] input, double[] expected) {
		try (Arena a = Arena.ofConfined()) {
			MemorySegment n = a.allocate(java.lang.foreign.ValueLayout.JAVA_LONG);
			B2nn.check((int) B2nn.NUM_WEIGHTS.invokeExact(handle, n));
			long count = n.get(java.lang.foreign.ValueLayout.JAVA_LONG, 0);
			MemorySegment g = a.allocate(JAVA_DOUBLE, count);
			B2nn.check((int) B2nn.WEIGHT_GRADIENT.invokeExact(handle, B2nn.pack(a, input), B2nn.pack(a, expected), g));
			Map<Integer, double[][]> out = new HashMap<>();
			long off = 0;
			for (int l = 0; l + 1 < layerDims.length; l++) {
				double[][] m = new double[layerDims[l]][layerDims[l + 1]];
				for (int i = 0; i < layerDims[l]; i++) {
					MemorySegment.copy(g, JAVA_DOUBLE, off * Double.BYTES, m[i], 0, layerDims[l + 1]);
					off += layerDims[l + 1];
				}
				out.put(l, m);
			}
			return out;
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
	}

	/**
	 * Drop-in path: one native call per step.  The batch map is packed in its own iteration order,
	 * which is the order the reference sums it in (SoftmaxCrossEntropyNeuralNet.java:305).
	 */
	@Override
	public void gradientStep(Map<double[], double[]> batch, double step, double momentum, boolean noise) {
		if (batch == null || batch.isEmpty()) throw new IllegalArgumentException("empty batch");
		int B = batch.size(), d0 = layerDims[0], dL = getOutputDim();
		double[][] xs = new double[B][], ts = new double[B][];
		int r = 0;
		for (Map.Entry<double[], double[]> e : batch.entrySet()) {
			xs[r] = e.getKey();
			ts[r] = e.getValue();
			r++;
		}
		try (Arena a = Arena.ofConfined()) {
			B2nn.check((int) B2nn.GRADIENT_STEP.invokeExact(handle, B2nn.pack(a, xs, d0), B2nn.pack(a, ts, dL), B,
					step, momentum, noise ? 1 : 0, MemorySegment.NULL));
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
	}

	@Override
	public int getInputDim() {
		return layerDims[0];
	}

	@Override
	public int getOutputDim() {
		return layerDims[layerDims.length - 1];
	}

	/**
	 * Not in the reference (it has no persistence, NeuralNet.java:7-67): weights, previousUpdate and time
	 * (SoftmaxCrossEntropyNeuralNet.java:327-344) to one flat file, so a long MNISTTrainer run can stop and resume.
	 */
	public void saveCheckpoint(String path) {
		try (Arena a = Arena.ofConfined()) {
			B2nn.check((int) B2nn.CHECKPOINT_SAVE.invokeExact(handle, a.allocateFrom(path)));
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
	}

	/** restores what saveCheckpoint wrote; IllegalArgumentException when layerDims, class or functions differ */
	public void loadCheckpoint(String path) {
		try (Arena a = Arena.ofConfined()) {
			B2nn.check((int) B2nn.CHECKPOINT_LOAD.invokeExact(handle, a.allocateFrom(path)));
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
	}

	/** `time` (SoftmaxCrossEntropyNeuralNet.java:23): gradient steps taken */
	public long getTime() {
		try (Arena a = Arena.ofConfined()) {
			MemorySegment steps = a.allocate(java.lang.foreign.ValueLayout.JAVA_LONG);
			B2nn.check((int) B2nn.TIME.invokeExact(handle, steps));
			return steps.get(java.lang.foreign.ValueLayout.JAVA_LONG, 0);
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
	}

	static int[] checkedDims(MemorySegment handle) throws Throwable {
		try (Arena a = Arena.ofConfined()) {
			MemorySegment in = a.allocate(JAVA_INT), out = a.allocate(JAVA_INT);
			B2nn.check((int) B2nn.DIMS.invokeExact(handle, in, out));
			return new int[]{in.get(JAVA_INT, 0), out.get(JAVA_INT, 0)};
		}
	}
}
