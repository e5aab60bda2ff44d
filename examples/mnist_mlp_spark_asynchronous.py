"""Hogwild training against the device-resident parameter server, pushing after every batch
(reference: examples/mnist_mlp_spark_asynchronous.py; `parameter_server_mode='nvlink'` is the B200 addition)."""
from _data import mnist_like
from elephas.spark_model import SparkModel
from elephas.utils.rdd_utils import to_simple_rdd
from elephas_b200.models import SGD, zoo
from elephas_b200.spark import SparkConf, SparkContext

sc = SparkContext(conf=SparkConf().setAppName('Mnist_Spark_MLP').setMaster('local[4]'))
x_train, y_train, x_test, y_test = mnist_like()
model = zoo.mnist_mlp()
model.compile(SGD(learning_rate=0.05), 'categorical_crossentropy', ['acc'])
spark_model = SparkModel(model, frequency='batch', mode='hogwild', parameter_server_mode='nvlink', num_workers=4)
spark_model.fit(to_simple_rdd(sc, x_train, y_train), epochs=1, batch_size=64, verbose=0, validation_split=0.1)
print('Test accuracy:', spark_model.evaluate(x_test, y_test)[1])
