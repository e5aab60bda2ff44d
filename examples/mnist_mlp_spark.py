"""MNIST-shaped MLP, asynchronous mode (reference: examples/mnist_mlp_spark.py).
Run: python examples/mnist_mlp_spark.py            (one device)
     python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/mnist_mlp_spark.py"""
from _data import mnist_like
from elephas.spark_model import SparkModel
from elephas.utils.rdd_utils import to_simple_rdd
from elephas_b200.models import SGD, Activation, Dense, Dropout, Sequential
from elephas_b200.spark import SparkConf, SparkContext

batch_size, epochs = 64, 1
conf = SparkConf().setAppName('Mnist_Spark_MLP').setMaster('local[8]')
sc = SparkContext(conf=conf)
x_train, y_train, x_test, y_test = mnist_like()
print(x_train.shape[0], 'train samples', x_test.shape[0], 'test samples')

model = Sequential()
model.add(Dense(128, input_dim=784))
model.add(Activation('relu'))
model.add(Dropout(0.2))
model.add(Dense(128))
model.add(Activation('relu'))
model.add(Dropout(0.2))
model.add(Dense(10))
model.add(Activation('softmax'))
model.compile(SGD(learning_rate=0.1), 'categorical_crossentropy', ['acc'])

rdd = to_simple_rdd(sc, x_train, y_train)
spark_model = SparkModel(model, frequency='epoch', mode='asynchronous')
spark_model.fit(rdd, epochs=epochs, batch_size=batch_size, verbose=0, validation_split=0.1)
score = spark_model.evaluate(x_test, y_test, verbose=2)
print('Test accuracy:', score[1])
