import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.SymbolLookup;
import java.lang.foreign.ValueLayout;
import java.lang.invoke.MethodHandle;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

/**
 * Panama FFM (JDK 22+) downcall handles for libb2nn.so (include/b2nn.h).  One static handle per
 * export; every wrapper turns a negative status into the Java exception the reference's
 * preconditions would have raised (assert -> IllegalArgumentException, etc.).
 *
 * NOT COMPILED IN THIS REPOSITORY'S CI: the build and GPU boxes carry no JDK (DESIGN.md 1).
 * Default package on purpose -- the reference lives in the default package.
 */
final class B2nn {
	static final int OK = 0, ERR_INVALID_ARG = -1, ERR_DIM_MISMATCH = -2, ERR_UNSUPPORTED = -3;
	static final int NET_SOFTMAX_CE = 0, NET_GENERAL = 1;
	static final int PREC_F64 = 0, PREC_BF16 = 1, PREC_TF32 = 2;
	static final int DT_F64 = 0;
	static final int DATA_DEVICE = 0;

	private static final Linker LINKER = Linker.nativeLinker();
	private static final SymbolLookup LIB = SymbolLookup.libraryLookup(
			System.getProperty("b2nn.lib", "libb2nn.so"), Arena.global());

	private static MethodHandle h(String name, FunctionDescriptor fd) {
		return LINKER.downcallHandle(LIB.find(name).orElseThrow(
				() -> new UnsatisfiedLinkError("libb2nn.so does not export " + name)), fd);
	}

	static final MethodHandle CREATE = h("b2nn_create", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT,
			JAVA_INT, JAVA_INT, JAVA_DOUBLE, JAVA_INT, JAVA_DOUBLE, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_LONG,
			JAVA_INT, ADDRESS));
	static final MethodHandle DESTROY = h("b2nn_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS));
	static final MethodHandle DIMS = h("b2nn_dims", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
	static final MethodHandle NUM_WEIGHTS = h("b2nn_num_weights", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
	static final MethodHandle PROPAGATE = h("b2nn_propagate",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, ADDRESS));
	static final MethodHandle LOSS = h("b2nn_loss",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_INT, ADDRESS));
	static final MethodHandle WEIGHT_GRADIENT = h("b2nn_weight_gradient",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
	static final MethodHandle GRADIENT_STEP = h("b2nn_gradient_step", FunctionDescriptor.of(JAVA_INT, ADDRESS,
			ADDRESS, ADDRESS, JAVA_INT, JAVA_DOUBLE, JAVA_DOUBLE, JAVA_INT, ADDRESS));
	static final MethodHandle DATASET_CREATE = h("b2nn_dataset_create", FunctionDescriptor.of(JAVA_INT, ADDRESS,
			ADDRESS, JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, JAVA_INT, ADDRESS));
	static final MethodHandle DATASET_DESTROY = h("b2nn_dataset_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS));
	static final MethodHandle TRAIN_INDEXED = h("b2nn_train_indexed", FunctionDescriptor.of(JAVA_INT, ADDRESS,
			ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, JAVA_DOUBLE, JAVA_DOUBLE, ADDRESS));
	static final MethodHandle VALIDATE = h("b2nn_validate",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, ADDRESS));
	static final MethodHandle EVALUATE = h("b2nn_evaluate",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS));
	static final MethodHandle SYNC = h("b2nn_sync", FunctionDescriptor.of(JAVA_INT, ADDRESS));
	static final MethodHandle TIME = h("b2nn_time", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
	static final MethodHandle CHECKPOINT_SAVE = h("b2nn_checkpoint_save",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
	static final MethodHandle CHECKPOINT_LOAD = h("b2nn_checkpoint_load",
			FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
	static final MethodHandle LAST_ERROR = h("b2nn_last_error", FunctionDescriptor.of(ADDRESS));

	private B2nn() {
	}

	static String lastError() {
		try {
			MemorySegment p = (MemorySegment) LAST_ERROR.invokeExact();
			return p.reinterpret(4096).getString(0);
		} catch (Throwable t) {
			return "(b2nn_last_error failed: " + t + ")";
		}
	}

	/** maps B2NN_ERR_* to the exception type the Java caller expects */
	static void check(int status) {
		if (status == OK) return;
		String msg = lastError();
		switch (status) {
			case ERR_INVALID_ARG:
			case ERR_DIM_MISMATCH:
				throw new IllegalArgumentException(msg);
			case ERR_UNSUPPORTED:
				throw new UnsupportedOperationException(msg);
			default:
				throw new IllegalStateException("b2nn status " + status + ": " + msg);
		}
	}

	/** copies double rows into one contiguous native segment, row-major */
	static MemorySegment pack(Arena arena, double[][] rows, int width) {
		MemorySegment seg = arena.allocate(JAVA_DOUBLE, (long) rows.length * width);
		for (int r = 0; r < rows.length; r++) {
			if (rows[r].length != width) throw new IllegalArgumentException("row " + r + " has the wrong dimension");
			MemorySegment.copy(rows[r], 0, seg, JAVA_DOUBLE, (long) r * width * Double.BYTES, width);
		}
		return seg;
	}

	static MemorySegment pack(Arena arena, double[] row) {
		return arena.allocateFrom(JAVA_DOUBLE, row);
	}

	static MemorySegment pack(Arena arena, int[] v) {
		return arena.allocateFrom(ValueLayout.JAVA_INT, v);
	}
}
