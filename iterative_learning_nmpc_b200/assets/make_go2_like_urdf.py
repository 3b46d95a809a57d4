"""Write `go2_like.urdf`: a SYNTHETIC Go2-like quadruped description.

The reference obtains the Go2 URDF from the `robot_descriptions` package through
`mj_pin.utils.get_robot_description` (`Behavior_Cloning/utils/Rollout_MPC.py:320`); neither the
package nor any URDF exists in this environment.  The numbers
below approximate the public Unitree Go2 description (12 revolute joints, hip-x / thigh-y /
calf-y, foot frames `FL_foot`..`RR_foot` attached by fixed joints).  They are NOT the real file;
a real `go2.urdf` of the same topology can be passed to the code generator instead.
"""
import os

LEGS = {  # name: (sign_x, sign_y)
    "FL": (+1, +1),
    "FR": (+1, -1),
    "RL": (-1, +1),
    "RR": (-1, -1),
}


def inertial(mass, com, diag, off=(0.0, 0.0, 0.0)):
    return (f'    <inertial>\n'
            f'      <origin xyz="{com[0]} {com[1]} {com[2]}" rpy="0 0 0"/>\n'
            f'      <mass value="{mass}"/>\n'
            f'      <inertia ixx="{diag[0]}" ixy="{off[0]}" ixz="{off[1]}" '
            f'iyy="{diag[1]}" iyz="{off[2]}" izz="{diag[2]}"/>\n'
            f'    </inertial>\n')


def main():
    out = ['<?xml version="1.0"?>\n',
           '<!-- SYNTHETIC Go2-like model, see make_go2_like_urdf.py.  Not the Unitree file. -->\n',
           '<robot name="go2">\n',
           '  <link name="base">\n',
           inertial(6.921, (0.021112, 0.0, -0.005366), (0.02448, 0.098077, 0.107),
                    (0.00012166, 0.0014849, -0.0000312)),
           '  </link>\n']
    for leg, (sx, sy) in LEGS.items():
        out += [
            f'  <joint name="{leg}_hip_joint" type="revolute">\n'
            f'    <origin xyz="{sx * 0.1934} {sy * 0.0465} 0" rpy="0 0 0"/>\n'
            f'    <parent link="base"/>\n    <child link="{leg}_hip"/>\n'
            f'    <axis xyz="1 0 0"/>\n'
            f'    <limit lower="-1.0472" upper="1.0472" effort="23.7" velocity="30.1"/>\n'
            f'  </joint>\n',
            f'  <link name="{leg}_hip">\n',
            inertial(0.678, (sx * -0.0054, sy * 0.00194, -0.000105),
                     (0.00048, 0.000884, 0.000596),
                     (sx * sy * -0.00000301, sx * 0.00000111, sy * -0.00000142)),
            '  </link>\n',
            f'  <joint name="{leg}_thigh_joint" type="revolute">\n'
            f'    <origin xyz="0 {sy * 0.0955} 0" rpy="0 0 0"/>\n'
            f'    <parent link="{leg}_hip"/>\n    <child link="{leg}_thigh"/>\n'
            f'    <axis xyz="0 1 0"/>\n'
            f'    <limit lower="-1.5708" upper="3.4907" effort="23.7" velocity="30.1"/>\n'
            f'  </joint>\n',
            f'  <link name="{leg}_thigh">\n',
            inertial(1.152, (-0.00374, sy * -0.0223, -0.0327),
                     (0.00584, 0.0058, 0.00103),
                     (sy * 0.0000878, -0.000289, sy * 0.000808)),
            '  </link>\n',
            f'  <joint name="{leg}_calf_joint" type="revolute">\n'
            f'    <origin xyz="0 0 -0.213" rpy="0 0 0"/>\n'
            f'    <parent link="{leg}_thigh"/>\n    <child link="{leg}_calf"/>\n'
            f'    <axis xyz="0 1 0"/>\n'
            f'    <limit lower="-2.7227" upper="-0.83776" effort="45.43" velocity="15.7"/>\n'
            f'  </joint>\n',
            f'  <link name="{leg}_calf">\n',
            inertial(0.154, (0.00548, sy * -0.000975, -0.115),
                     (0.00108, 0.0011, 0.0000329),
                     (sy * 0.00000034, 0.0000172, sy * 0.00000835)),
            '  </link>\n',
            f'  <joint name="{leg}_foot_joint" type="fixed">\n'
            f'    <origin xyz="0 0 -0.213" rpy="0 0 0"/>\n'
            f'    <parent link="{leg}_calf"/>\n    <child link="{leg}_foot"/>\n'
            f'  </joint>\n',
            f'  <link name="{leg}_foot">\n',
            inertial(0.04, (0.0, 0.0, 0.0), (0.0000096, 0.0000096, 0.0000096)),
            '  </link>\n',
        ]
    out.append('</robot>\n')
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "go2_like.urdf")
    with open(path, "w") as f:
        f.writelines(out)
    print("wrote", path)


if __name__ == "__main__":
    main()
