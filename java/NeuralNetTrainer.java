import java.io.PrintStream;
import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.util.ArrayList;
import java.util.List;
import java.util.Map;
import java.util.Random;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_INT;

/**
 * Drop-in replacement for the reference's src/NeuralNetTrainer.java: same constructor (:28) and
 * train signature (:60).  dataMaster is uploaded to the GPU once (b2nn_dataset_create); sample()
 * keeps the reference's draw semantics (:143-158: nextInt(sampler.size()), get, remove, refill when
 * empty) but draws ROW NUMBERS, and whole runs of mini-batches go down in one native call
 * (b2nn_train_indexed) instead of one crossing per gradientStep.
 *
 * Master order = the iteration order of the map handed in (the reference's own order is the
 * identity-hash order of a HashMap, which is not reproducible; pass a LinkedHashMap for a fixed one).
 */
class NeuralNetTrainer {
	/** iterations shipped per native call when nothing has to be printed in between */
	private static final int CHUNK = 2000;

	private final B2nnNet net;
	private final int rows;
	private final MemorySegment dataset;
	private final List<Integer> dataSampler = new ArrayList<>();
	private final Random random = new Random(1);   // reference :42

	NeuralNetTrainer(Map<double[], double[]> data, NeuralNet net) {
		if (data == null || net == null) throw new IllegalArgumentException("data and net must be non-null");
		if (!(net instanceof B2nnNet)) {
			throw new IllegalArgumentException("this trainer drives the GPU backend only (no CPU fallback)");
		}
		this.net = (B2nnNet) net;
		this.rows = data.size();
		int d0 = net.getInputDim(), dL = net.getOutputDim();
		double[][] xs = new double[rows][], ts = new double[rows][];
		int r = 0;
		for (Map.Entry<double[], double[]> e : data.entrySet()) {
			xs[r] = e.getKey();
			ts[r] = e.getValue();
			r++;
		}
		try (Arena a = Arena.ofConfined()) {
			MemorySegment out = a.allocate(ADDRESS);
			B2nn.check((int) B2nn.DATASET_CREATE.invokeExact(this.net.handle, B2nn.pack(a, xs, d0), B2nn.DT_F64,
					B2nn.pack(a, ts, dL), MemorySegment.NULL, (long) rows, B2nn.DATA_DEVICE, out));
			dataset = out.get(ADDRESS, 0);
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
		refillSampler();
	}

	void train(int iterations, double stepSize, int batchSize, double momentum, boolean noise, ProgressBar monitor,
			PrintStream observer) {
		if (iterations <= 0 || stepSize <= 0) throw new IllegalArgumentException("iterations, stepSize > 0 (ref :62)");
		if (batchSize <= 0 || batchSize >= rows) throw new IllegalArgumentException("0 < batchSize < rows (ref :63)");
		if (noise) throw new UnsupportedOperationException("noise=true: the reference's term is NaN for negative draws");
		int validationSize = rows / 100 + 1;   // reference :65
		int chunk = observer != null ? 1 : CHUNK;
		try (Arena a = Arena.ofConfined()) {
			MemorySegment idx = a.allocate(JAVA_INT, (long) chunk * batchSize);
			MemorySegment loss = a.allocate(JAVA_DOUBLE);
			for (int done = 0; done < iterations; ) {
				int n = Math.min(chunk, iterations - done);
				for (int i = 0; i < n * batchSize; i++) idx.setAtIndex(JAVA_INT, i, sampleRow());
				B2nn.check((int) B2nn.TRAIN_INDEXED.invokeExact(net.handle, dataset, idx, n, batchSize, stepSize,
						momentum, MemorySegment.NULL));
				if (observer != null) {   // reference :71 -- "%d,%.2f\n" per iteration
					B2nn.check((int) B2nn.VALIDATE.invokeExact(net.handle, dataset, validationSize, loss));
					observer.printf("%d,%.2f\n", done, loss.get(JAVA_DOUBLE, 0));
				}
				if (monitor != null) for (int i = 0; i < n; i++) monitor.step();
				done += n;
			}
			B2nn.check((int) B2nn.SYNC.invokeExact(net.handle));
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
	}

	/** `>=`-argmax hit rate over the first n master rows (MNISTTrainer.java:180-197 on the training partition) */
	double accuracy(int n) {
		try (Arena a = Arena.ofConfined()) {
			MemorySegment hits = a.allocate(java.lang.foreign.ValueLayout.JAVA_LONG);
			B2nn.check((int) B2nn.EVALUATE.invokeExact(net.handle, dataset, (long) n, hits, MemorySegment.NULL,
					MemorySegment.NULL));
			return hits.get(java.lang.foreign.ValueLayout.JAVA_LONG, 0) / (double) n;
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
	}

	/** one draw of sample() (reference :151-154) on row numbers */
	private int sampleRow() {
		if (dataSampler.isEmpty()) refillSampler();
		int r = random.nextInt(dataSampler.size());
		int row = dataSampler.get(r);
		dataSampler.remove(r);
		return row;
	}

	private void refillSampler() {
		dataSampler.clear();
		for (int i = 0; i < rows; i++) dataSampler.add(i);
	}
}
