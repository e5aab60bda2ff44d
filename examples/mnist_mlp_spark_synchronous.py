"""MNIST-shaped MLP, synchronous mode, 2 workers (reference: examples/mnist_mlp_spark_synchronous.py)."""
from _data import mnist_like
from elephas.spark_model import SparkModel
from elephas.utils.rdd_utils import to_simple_rdd
from elephas_b200.models import SGD, Activation, Dense, Dropout, Sequential
from elephas_b200.spark import SparkConf, SparkContext

sc = SparkContext(conf=SparkConf().setAppName('Mnist_Spark_MLP').setMaster('local[8]'))
x_train, y_train, x_test, y_test = mnist_like()
model = Sequential([Dense(128, input_dim=784), Activation('relu'), Dropout(0.2), Dense(128), Activation('relu'),
                    Dropout(0.2), Dense(10), Activation('softmax')])
model.compile(SGD(learning_rate=0.1), 'categorical_crossentropy', ['acc'])
spark_model = SparkModel(model, mode='synchronous', num_workers=2)
spark_model.fit(to_simple_rdd(sc, x_train, y_train), epochs=2, batch_size=64, verbose=2, validation_split=0.1)
print('Test accuracy:', spark_model.evaluate(x_test, y_test, verbose=2)[1])
print('histories:', [h['loss'] for h in spark_model.training_histories])
