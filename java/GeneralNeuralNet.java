/**
 * Drop-in replacement for the reference's src/GeneralNeuralNet.java (constructor :112-115):
 * configurable inner / last activation and loss, selected from B2nnFunctions' singletons and
 * mapped to the kernel enums of include/b2nn.h.
 */
class GeneralNeuralNet extends B2nnNet {
	public GeneralNeuralNet(int[] layerDims,
			ActivationFunction innerActivationFunc, ActivationPrime innerActivationPrime,
			ActivationFunction lastActivationFunc, ActivationPrime lastActivationPrime,
			LossFunction lossFunc, LossFunctionPrime lossPrime) {
		super(layerDims, B2nn.NET_GENERAL,
				B2nnFunctions.require(innerActivationFunc, innerActivationPrime, "inner activation").kernelId(),
				B2nnFunctions.require(innerActivationFunc, innerActivationPrime, "inner activation").kernelParam(),
				B2nnFunctions.require(lastActivationFunc, lastActivationPrime, "last activation").kernelId(),
				B2nnFunctions.require(lastActivationFunc, lastActivationPrime, "last activation").kernelParam(),
				B2nnFunctions.require(lossFunc, lossPrime, "loss").kernelId());
	}
}
