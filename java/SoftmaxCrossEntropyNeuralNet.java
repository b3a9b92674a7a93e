/**
 * Drop-in replacement for the reference's src/SoftmaxCrossEntropyNeuralNet.java: same class name,
 * package, visibility and constructor (reference :103); every method body (propagate :164,
 * calculateLoss :207, calculateWeightGradient :229, gradientStep :297) now runs on the GPU
 * through libb2nn.so (see B2nnNet).  The activation pair must be one of B2nnFunctions'
 * singletons.
 */
public class SoftmaxCrossEntropyNeuralNet extends B2nnNet {
	SoftmaxCrossEntropyNeuralNet(int[] layerDims, ActivationFunction innerActivationFunc,
			ActivationPrime innerActivationPrime) {
		super(layerDims, B2nn.NET_SOFTMAX_CE,
				B2nnFunctions.require(innerActivationFunc, innerActivationPrime, "inner activation").kernelId(),
				B2nnFunctions.require(innerActivationFunc, innerActivationPrime, "inner activation").kernelParam(),
				0, 0.0, 0);
	}
}
