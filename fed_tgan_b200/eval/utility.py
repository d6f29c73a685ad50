"""ML utility: train LR / DT / RF / MLP on real vs. synthetic, test on real.

Parity: ``Server/utility_analysis.py:15-119`` (label-encode on train+test, StandardScaler
fitted on train+test features, ``class_weight='balanced', random_state=69``; reports the
matrix of (accuracy, weighted-F1) differences and ``difference in f1-score`` = mean F1 gap).
"""

from __future__ import annotations

import argparse

import numpy as np
import pandas as pd
from sklearn import ensemble, linear_model, metrics, preprocessing, tree
from sklearn.neural_network import MLPClassifier

from ..data.schemas import INTRUSION_CATEGORICAL


def real_res(original_real, real, test, target_col, cat_cols, verbose=True):
    real, test, original_real = real.copy(), test.copy(), original_real.copy()
    for c in cat_cols or []:
        le = preprocessing.LabelEncoder()
        for frame in (real, test, original_real):
            frame[c] = frame[c].astype(str)
        le.fit(original_real[c].values)
        real[c], test[c], original_real[c] = le.transform(real[c]), le.transform(test[c]), le.transform(original_real[c])
    y_tr, x_tr = real[target_col], real.drop(columns=[target_col])
    y_te, x_te = test[target_col], test.drop(columns=[target_col])
    scaler = preprocessing.StandardScaler().fit(original_real.drop(columns=[target_col]).values)
    x_tr, x_te = scaler.transform(x_tr.values), scaler.transform(x_te.values)
    models = [
        ("LR", linear_model.LogisticRegression(class_weight="balanced", random_state=69)),
        ("DT", tree.DecisionTreeClassifier(class_weight="balanced", random_state=69)),
        ("RF", ensemble.RandomForestClassifier(class_weight="balanced", random_state=69)),
        ("MLP", MLPClassifier(random_state=69)),
    ]
    out = []
    for name, m in models:
        if verbose:
            print(f"training of {name}")
        pred = m.fit(x_tr, y_tr).predict(x_te)
        out.append([metrics.accuracy_score(y_te, pred), metrics.f1_score(y_te, pred, average="weighted")])
    return out


def utility_difference(train_path, test_path, synthetic_path, cat_cols, target_col, verbose=True):
    real, test, fake = (pd.read_csv(p) if isinstance(p, str) else p for p in (train_path, test_path, synthetic_path))
    original = pd.concat([real, test])
    r = real_res(original, real, test, target_col, cat_cols, verbose)
    f = real_res(original, fake, test, target_col, cat_cols, verbose)
    diff = np.asarray(r) - np.asarray(f)
    return diff, float(diff.mean(axis=0)[1])


def main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("-train_path", default="data/raw/Intrusion_train.csv")
    p.add_argument("-test_path", default="data/raw/Intrusion_test.csv")
    p.add_argument("-synthetic_path", default="Intrusion_result/Intrusion_synthesis_epoch_0.csv")
    p.add_argument("-target_column", default="class")
    p.add_argument("-categorical_list", default=None)
    a = p.parse_args(argv)
    cats = a.categorical_list.split(",") if a.categorical_list else list(INTRUSION_CATEGORICAL)
    diff, f1 = utility_difference(a.train_path, a.test_path, a.synthetic_path, cats, a.target_column)
    print("difference in accuracy and f1-score for all AL algorithms: ", diff)
    print("difference in f1-score: ", f1)
    return f1
