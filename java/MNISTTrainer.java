import java.io.DataInputStream;
import java.io.File;
import java.io.FileInputStream;
import java.io.FileNotFoundException;
import java.io.FileOutputStream;
import java.io.IOException;
import java.io.BufferedInputStream;
import java.io.PrintStream;
import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.util.LinkedHashMap;
import java.util.Map;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

/**
 * Drop-in replacement for the reference's src/MNISTTrainer.java on the GPU backend: same entry point
 * ({@code main}, reference :221-245, with its literals), same data files (reference :28-31), same CSV row
 * (reference :211-219), same evaluation order (training accuracy first, :241).  What differs:
 * <ul>
 * <li>the two inline leaky-ReLU lambdas of the reference's main (:234-235) become the
 *     {@link B2nnFunctions#LEAKY_RELU} singletons -- the backend maps the built-in functions to kernel enums and has
 *     no CPU fallback for arbitrary lambdas;</li>
 * <li>testOnTrainingData / testOnTestData (reference :159-197: one propagate call per row, 70 000 per run) are ONE
 *     native call each: b2nn_evaluate runs the forward pass and the {@code >=} argmax (last maximum wins, :166-168)
 *     over the whole partition on the device;</li>
 * <li>the partitions are kept in file order (LinkedHashMap / arrays) instead of identity-hash order, so a run is
 *     reproducible; the IDX magic numbers are checked unconditionally (the reference checks them inside
 *     {@code assert}, which silently shifts the parse when assertions are off).</li>
 * </ul>
 * Not compiled on the build box (no JDK there); written against JDK 22 java.lang.foreign like the rest of java/.
 */
class MNISTTrainer {
	private final B2nnNet net;
	private final NeuralNetTrainer trainer;
	private final int trainingRows;
	private final double[][] testImages;
	private final int[] testLabels;
	private final PrintStream trainingLog;

	private MNISTTrainer(NeuralNet net) throws IOException {
		if (!(net instanceof B2nnNet)) throw new IllegalArgumentException("GPU backend nets only (no CPU fallback)");
		this.net = (B2nnNet) net;
		Map<double[], double[]> trainingPartition = new LinkedHashMap<>();
		try (DataInputStream labels = open("data/train-labels-idx1-ubyte");
				DataInputStream images = open("data/train-images-idx3-ubyte")) {
			int n = header(labels, images);
			int bytes = images.readInt() * images.readInt();
			if (bytes != net.getInputDim()) throw new IOException("image size " + bytes + " != net input dimension");
			for (int i = 0; i < n; i++) {
				double[] oneHot = new double[10];
				oneHot[labels.readUnsignedByte()] = 1.0;
				trainingPartition.put(pixels(images, bytes), oneHot);
			}
		}
		try (DataInputStream labels = open("data/t10k-labels-idx1-ubyte");
				DataInputStream images = open("data/t10k-images-idx3-ubyte")) {
			int n = header(labels, images);
			int bytes = images.readInt() * images.readInt();
			if (bytes != net.getInputDim()) throw new IOException("test image size differs from the training set's");
			testImages = new double[n][];
			testLabels = new int[n];
			for (int i = 0; i < n; i++) {
				testImages[i] = pixels(images, bytes);
				testLabels[i] = labels.readUnsignedByte();
			}
		}
		trainingLog = new PrintStream(new FileOutputStream("logs/trainLog.csv", true));   // append, opened before training
		trainingRows = trainingPartition.size();
		trainer = new NeuralNetTrainer(trainingPartition, net);
	}

	private static DataInputStream open(String path) throws FileNotFoundException {
		return new DataInputStream(new BufferedInputStream(new FileInputStream(new File(path)), 1 << 20));
	}

	/** magic numbers 2049 / 2051 and the two sample counts, which must agree */
	private static int header(DataInputStream labels, DataInputStream images) throws IOException {
		if (labels.readInt() != 2049 || images.readInt() != 2051) throw new IOException("not an IDX1 / IDX3 file pair");
		int n = labels.readInt();
		if (images.readInt() != n) throw new IOException("label and image counts differ");
		return n;
	}

	/** one image as doubles in [0, 1]: unsigned byte / 255.0, exactly the reference's normalisation (:97) */
	private static double[] pixels(DataInputStream in, int bytes) throws IOException {
		byte[] raw = new byte[bytes];
		in.readFully(raw);
		double[] out = new double[bytes];
		for (int i = 0; i < bytes; i++) out[i] = (raw[i] & 0xff) / 255.0;
		return out;
	}

	private void train(int iterations, double stepSize, int batchSize, double momentum, boolean noise,
			boolean monitoring, boolean logging) throws FileNotFoundException {
		ProgressBar pb = monitoring ? new ProgressBar(10, iterations, progress -> "▯") : null;
		PrintStream log = null;
		if (logging) {
			long id = System.nanoTime() % 99999;
			log = new PrintStream(new File("logs/log" + id + ".csv"));
			System.out.println("Logging net performance to " + id);
		}
		trainer.train(iterations, stepSize, batchSize, momentum, noise, pb, log);
		if (log != null) log.close();
	}

	/** hit rate over the test partition: one b2nn_evaluate call over a labelled device data set */
	private double testOnTestData() {
		try (Arena a = Arena.ofConfined()) {
			int n = testImages.length;
			MemorySegment out = a.allocate(ADDRESS), hits = a.allocate(JAVA_LONG);
			B2nn.check((int) B2nn.DATASET_CREATE.invokeExact(net.handle, B2nn.pack(a, testImages, testImages[0].length),
					B2nn.DT_F64, MemorySegment.NULL, B2nn.pack(a, testLabels), (long) n, B2nn.DATA_DEVICE, out));
			MemorySegment ds = out.get(ADDRESS, 0);
			try {
				B2nn.check((int) B2nn.EVALUATE.invokeExact(net.handle, ds, (long) n, hits, MemorySegment.NULL,
						MemorySegment.NULL));
			} finally {
				int ignored = (int) B2nn.DATASET_DESTROY.invokeExact(ds);
			}
			return hits.get(JAVA_LONG, 0) / (double) n;
		} catch (RuntimeException | Error e) {
			throw e;
		} catch (Throwable t) {
			throw new IllegalStateException(t);
		}
	}

	/** hit rate over the training partition: the trainer's device-resident dataMaster, expected class = the last index
	 *  whose target entry equals 1 (reference :183-186), derived by the library when the data set is staged */
	private double testOnTrainingData() {
		return trainer.accuracy(trainingRows);
	}

	private void logTest(int[] netDim, int iterations, double stepSize, int batchSize, double momentum, boolean noise,
			double trainingAcc, double testAcc) {
		StringBuilder dims = new StringBuilder().append(netDim[0]);
		for (int i = 1; i < netDim.length; i++) dims.append('-').append(netDim[i]);
		trainingLog.printf("%s,%d,%.5f,%d,%.3f,%s,%.5f,%.5f\n", dims, iterations, stepSize, batchSize, momentum, noise,
				trainingAcc, testAcc);
		trainingLog.flush();
	}

	public static void main(String[] args) {
		try {
			int[] netDim = new int[]{784, 300, 100, 10};
			int iterations = 100000;
			double stepSize = 0.0125;
			int batchSize = 16;
			double momentum = 0.9;
			boolean noise = false;

			// the ONE edit of the reference's main: the built-in function objects instead of inline lambdas
			NeuralNet net = new SoftmaxCrossEntropyNeuralNet(netDim, B2nnFunctions.LEAKY_RELU,
					B2nnFunctions.LEAKY_RELU_PRIME);
			MNISTTrainer trainer = new MNISTTrainer(net);
			trainer.train(iterations, stepSize, batchSize, momentum, noise, true, false);
			double trainingAcc = trainer.testOnTrainingData();   // evaluation order of the reference: training set first
			double testAcc = trainer.testOnTestData();
			trainer.logTest(netDim, iterations, stepSize, batchSize, momentum, noise, trainingAcc, testAcc);
		} catch (IOException e) {
			e.printStackTrace();
		}
	}
}
