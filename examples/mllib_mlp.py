"""Training from an RDD of MLlib LabeledPoints (reference: examples/mllib_mlp.py)."""
from _data import mnist_like
from elephas.spark_model import SparkMLlibModel
from elephas.utils.rdd_utils import to_labeled_point
from elephas_b200.models import RMSprop, zoo
from elephas_b200.spark import SparkConf, SparkContext

x_train, y_train, x_test, y_test = mnist_like(3000, 500)
model = zoo.mnist_mlp()
model.compile(RMSprop(), "categorical_crossentropy", ['acc'])
sc = SparkContext(conf=SparkConf().setAppName('Mnist_Spark_MLP').setMaster('local[8]'))
lp_rdd = to_labeled_point(sc, x_train, y_train, categorical=True)
spark_model = SparkMLlibModel(model=model, frequency='epoch', mode='synchronous')
spark_model.fit(lp_rdd, epochs=5, batch_size=32, verbose=0, validation_split=0.1, categorical=True, nb_classes=10)
print('Test accuracy:', spark_model.master_network.evaluate(x_test, y_test, verbose=2)[1])
