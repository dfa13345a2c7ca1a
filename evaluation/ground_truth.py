"""Ground-truth algorithm: an offline multinomial logistic regression trained on the WHOLE training set
for several epochs, scored on the test set — what the reference does with datawig's SimpleImputer in
python-ground-truth-algorithm.ipynb (weighted F1 0.47 on the real data, README.md:221-233).
Prints a classification report and the trainer's samples/s.  Usage: ground_truth.py [train.csv test.csv]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sklearn.linear_model import SGDClassifier
from sklearn.metrics import classification_report

from psb200.data import ReviewGenerator, read_csv


def main():
    if len(sys.argv) >= 3:
        train, test = read_csv(sys.argv[1]), read_csv(sys.argv[2])
    else:
        gen = ReviewGenerator(mean_len=420.0)
        train, test = gen.train_split(12000), gen.test_split()
    Xtr, Xte = train.to_dense(), test.to_dense()
    clf = SGDClassifier(loss="log_loss", alpha=1e-6, learning_rate="constant", eta0=0.5, random_state=0)
    classes = np.unique(train.labels)
    for epoch in range(10):
        t0 = time.time()
        order = np.random.default_rng(epoch).permutation(len(train))
        for i in range(0, len(order), 16):
            b = order[i:i + 16]
            clf.partial_fit(Xtr[b], train.labels[b], classes=classes)
        dt = time.time() - t0
        acc = (clf.predict(Xte) == test.labels).mean()
        print(f"Epoch[{epoch}] Speed: {len(train) / dt:.2f} samples/sec  Validation-Score-accuracy={acc:.6f}")
    print(classification_report(test.labels, clf.predict(Xte)))


if __name__ == "__main__":
    main()
